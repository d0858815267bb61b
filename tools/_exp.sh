mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "update_without_readback" 2>&1 | tail -15
