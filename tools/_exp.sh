mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "separable or conv1d" 2>&1 | tail -25
