"""Host-side checks of the synthetic workload generators: determinism by global index and, for
W4, that the emitted Dusp text and the directly built unit graph are the same circuit."""
import numpy as np

import dusp_b200 as d
from dusp_b200 import capi, config, workloads as W
from dusp_b200.config import _mix64
from dusp_b200.render import circuit_blob


def test_generators_are_functions_of_the_global_voice_index():
    a, b = W.w3_params(0, 8), W.w3_params(4, 4)
    for k in a:
        assert np.array_equal(a[k][4:], b[k])
    assert np.array_equal(W.w2_params(0, 0, [3, 9])["f0"], W.w2_params(0, 10)["f0"][[3, 9]])
    even, odd = W.w5_split(5, 6)
    assert even == [6, 8, 10] and odd == [5, 7, 9]


def test_w4_text_and_graph_are_the_same_circuit(lib, oracle):
    units = []
    for i in range(12):
        unit, text = W.w4_circuit(i, with_text=True)
        blob, _ = circuit_blob(unit)
        want = oracle.render(blob, 0.03)
        config.seed(_mix64(W.W4_SEED * 0x20003 + i))  # same Noise stream seeds as the generator handed out
        blob_t, _ = circuit_blob(d.unDusp(text))
        got = oracle.render(blob_t, 0.03)
        assert len(got) == len(want) == 1 and np.array_equal(got[0], want[0]), text
        p = capi.Program(blob)
        units.append(p.num_units)
        assert 1 <= p.num_units <= W.W4_MAX_UNITS + 1
    assert len(set(units)) > 4  # the circuits really differ in shape
    a, b = (capi.Program(circuit_blob(W.w4_circuit(3))[0]) for _ in range(2))  # deterministic in the index
    assert a.disassemble() == b.disassemble() and a.params() == b.params()
