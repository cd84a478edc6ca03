# GPU call 26: state block (checkpoint / resume) test
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "state_block or set_params" > gpurun_out/c26_pytest.log 2>&1; tail -15 gpurun_out/c26_pytest.log
