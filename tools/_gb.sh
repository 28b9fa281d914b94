for dbg in 1 2; do
echo "== debug $dbg"
MDX_GEMM_DEBUG=$dbg python tools/gemm_bench.py 524288 2560 320 --geglu
MDX_GEMM_DEBUG=$dbg python tools/gemm_bench.py 524288 2560 320 --geglu --bn 128
MDX_GEMM_DEBUG=$dbg python tools/gemm_bench.py 524288 320 320 --out f32 --res f32
MDX_GEMM_DEBUG=$dbg python tools/gemm_bench.py 524288 960 320
done
