import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


# a small UNet that exercises every layer kind (ResBlock with/without skip conv, SpatialTransformer at three
# resolutions, Downsample, Upsample, skip concat) at channel counts the tensor-core path accepts (multiples of 64)
TINY_UNET = dict(in_channels=4, out_channels=4, model_channels=64, attention_resolutions=[4, 2, 1],
                 num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_head_channels=64, num_heads=-1,
                 use_spatial_transformer=True, use_linear_in_transformer=True, transformer_depth=1,
                 context_dim=128, legacy=False)
# SD-v1 / wukong shape: fixed 8 heads (d_head 8/16/32 here; 40/80/160 at full size), conv1x1 projections
TINY_UNET_V1 = dict(in_channels=4, out_channels=4, model_channels=64, attention_resolutions=[4, 2, 1],
                    num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_head_channels=-1, num_heads=8,
                    use_spatial_transformer=True, use_linear_in_transformer=False, transformer_depth=1,
                    context_dim=192, legacy=False)


@pytest.fixture(scope="session")
def tiny_cfg():
    return dict(TINY_UNET)


@pytest.fixture(scope="session")
def tiny_v1_cfg():
    return dict(TINY_UNET_V1)


def parity_log(name, value, tol):
    """Append a measured parity number to gpurun_out/parity_errors.jsonl (scratch; summarised under profiles/), so the
    errors quoted in BASELINE.md are the ones the tests asserted."""
    import json
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_errors.jsonl"), "a") as f:
            f.write(json.dumps({"test": name, "value": float(value), "tol": float(tol)}) + "\n")
    except OSError:
        pass
