# GPU call 34: interpreter instantiations now come from three translation units: the interpreter-mode parity tests, tile shapes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "interpreter or tile_shapes or w4 or multi" > gpurun_out/c34_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c34_pytest.log; tail -4 gpurun_out/c34_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
