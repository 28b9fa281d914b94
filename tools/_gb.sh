python tools/gemm_bench.py 524288 960 320
python tools/gemm_bench.py 524288 320 320 --out f32 --res f32
python tools/gemm_bench.py 524288 320 320 --out f32
python tools/gemm_bench.py 524288 320 320
python tools/gemm_bench.py 524288 2560 320 --geglu
python tools/gemm_bench.py 524288 2560 320 --geglu --bn 128
python tools/gemm_bench.py 131072 5120 640 --geglu
python tools/gemm_bench.py 32768 10240 1280 --geglu
python tools/gemm_bench.py 524288 320 1280 --res f32
python tools/gemm_bench.py 131072 1920 640
python tools/gemm_bench.py 131072 640 640 --out f32 --res f32
python tools/gemm_bench.py 524288 320 320 --taps 9 --hw 64
python tools/gemm_bench.py 524288 320 640 --taps 9 --hw 64 --out f32
python tools/gemm_bench.py 32768 1280 1280 --taps 9 --hw 16
python tools/gemm_bench.py 131072 640 1280 --taps 9 --hw 32
