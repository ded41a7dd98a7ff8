mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mnn_backend.py tests/test_gpu_parity.py -m gpu -x -q -k "dsp or mnn or plugin or identical" > gpurun_out/pytest_dsp.log 2>&1; echo rc=$?; tail -30 gpurun_out/pytest_dsp.log
