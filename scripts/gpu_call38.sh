# GPU call 38: the end-of-round build once more: full -m gpu suite, smoke()
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c38_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c38_pytest.log; tail -4 gpurun_out/c38_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
