"""C-ABI checks that need no GPU: the library loads, exports every declared
symbol, compiles circuits to programs and rejects what it does not accelerate."""
import ctypes
import re
import os

import pytest

import dusp_b200 as d
from dusp_b200 import capi, components as C, patches as P, quick
from dusp_b200.render import circuit_blob

import circuits

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_are_exported(lib):
    header = open(os.path.join(ROOT, "include", "dusp_b200.h")).read()
    declared = set(re.findall(r"\b(dusp_[a-z0-9_]+)\s*\(", header))
    assert declared == set(capi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.dusp_abi_version() == 1


def test_tick_generator_agrees_across_host_library_and_oracle(lib, oracle):
    for seed, k in [(0, 0), (77, 5), (2**64 - 1, 123456789), (0xDEADBEEF, 2**40), (0x0123456789ABCDEF, 2**32 - 1)]:
        assert lib.dusp_rand(seed, k) == d.config.rand_tick(seed, k) == oracle.rand(seed, k)
    u = [d.config.rand_tick(0xC0FFEE1234, k) for k in range(4096)]
    assert 0.0 <= min(u) and max(u) < 1.0 and abs(sum(u) / len(u) - 0.5) < 0.02 and len(set(u)) == len(u)


@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_every_case_compiles(lib, oracle, name):
    fac, _, _ = circuits.CASES[name]
    blob, circuit = circuit_blob(fac())
    # one program per epoch (graph edits; warm-up ticks while channel counts still grow): the library
    # splits the blob and compiles every section when the run is created (no GPU needed until the first render)
    run = capi.Run([blob], [1])
    epochs = run.epochs(0)
    assert [k for k, _ in epochs] == sorted(k for k, _ in epochs) and epochs[0][0] == 0
    for k, p in epochs:
        assert p.num_instructions >= 1 and 1 <= p.num_slots <= 32
        assert "out" in p.disassemble()
    assert p.num_units == len(circuit.units)  # the host circuit is left as the last epoch found it
    assert epochs[0][1].num_channels == len(oracle.render(blob, 0.005))  # one chunk
    assert run.num_channels == max(p.num_channels for _, p in epochs)
    run.close()


@pytest.mark.parametrize("name", ["w2_live", "w3_voice", "simple_delay", "transcendental"])
def test_specialised_kernel_compiles_for_sm100a(lib, name):
    """fused.cc: the circuit shape becomes one straight-line kernel built from the cores in ops.cuh;
    NVRTC targets sm_100a and needs no GPU."""
    blob, _ = circuit_blob(circuits.CASES[name][0]())
    p = capi.Program(blob)
    size, src = p.specialise(8, 4)
    assert size > 1000 and 'extern "C" __global__' in src and "#include \"ops.cuh\"" in src
    assert src.count("{ // ") == p.num_instructions


def test_program_parameters_are_named(lib):
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    names = dict(p.params())
    pan = [n for n in names if n.endswith(".pan[0]")]
    off = [n for n in names if n.endswith(".offset[0]")]
    seed = [n for n in names if n.endswith(".seed_lo")]
    assert len(pan) == 1 and len(off) == 3 and len(seed) == 1
    assert names[pan[0]] == pytest.approx(-0.3)
    assert p.bytes_per_voice_sample(capi.OUT_VOICES) == 24  # 8 out + 4 + 4 + 8
    assert p.bytes_per_voice_sample(capi.OUT_MIX) == 16


def test_cyclic_graph_uses_a_stale_buffer(lib):
    blob, _ = circuit_blob(circuits.simple_delay())
    text = capi.Program(blob).disassemble()
    assert "ldstale" in text and "ststale" in text and "1 stale" in text


def test_bad_blob_is_rejected(lib):
    with pytest.raises(d.DuspError):
        capi.Program("hello")
    with pytest.raises(d.DuspError, match="expects an outlet"):
        capi.Program("DUSPGRAPH 1\nEND\n")


def test_unsupported_configurations_fail_loudly(lib):
    # CircleBuffer taps closer than one chunk: sub-block order would differ from the reference, so the program is
    # marked for whole-chunk tiles instead (rendered, not refused: tests/circuits.py tight_taps)
    mt = P.MultiTapDelay(1, 1.0, C.Osc(100))
    tap = mt.addFeedback(0.001, 0.5, 0)
    blob, _ = circuit_blob(C.Sum(mt.writer.IN.outlet, tap))
    assert capi.Program(blob).num_instructions >= 5
    # a unit feeding itself
    m = C.Multiply(1, 0.5)
    m.A = m
    blob, _ = circuit_blob(m)
    with pytest.raises(capi.UnsupportedCircuit, match="own inlet"):
        capi.Program(blob)
    # events after the first chunk are carried as state assignments (EV records) ...
    blob, circ = circuit_blob(d.unDusp("[Osc f:([Ramp 100 1 2] at 1)]"))
    assert circ.device_events == [(172, 0, "playing", 1.0), (172, 0, "t", 0.0)] and "EV 172 0 playing 1" in blob
    # ... an event that rewires the graph starts a new EPOCH section holding the circuit after the edit ...
    from dusp_b200.render import split_epochs
    o = C.Osc(100)
    o.schedule(1.0, lambda unit: setattr(unit, "F", C.Osc(3)))
    blob, circ = circuit_blob(o)
    epochs = split_epochs(blob)
    assert [k for k, _ in epochs] == [0, 172] and len(circ.units) == 2
    assert capi.Program(epochs[0][1]).num_units == 1 and capi.Program(epochs[1][1]).num_units == 2
    # ... and an event that assigns anything but numbers and inlets is outside the accelerated path
    o = C.Osc(100)
    o.schedule(1.0, lambda unit: setattr(unit, "waveform", "saw"))
    with pytest.raises(capi.UnsupportedCircuit, match="event"):
        circuit_blob(o)
    # a Shape finishes from inside its own tick (Shape/index.js:40-42).  With a constant duration the sample is known: the
    # edit its onFinish makes becomes an epoch at the chunk that holds it, marked FIN (the oracle switches when ITS Shape
    # runs out) ...
    blob, _ = circuit_blob(d.unDusp("D0.1 then 0.5"))
    assert "EPOCH 17 FIN 0" in blob  # 0.1 s = 4410 samples: chunk 17
    # ... a signal-rate duration makes the moment data-dependent: refused; so is an onFinish that reaches a unit which
    # has already ticked in that tick; one that only raises finish flags (Boop.js:15-17) needs nothing
    env = C.Shape("decay", C.Sum(C.Multiply(C.Osc(1), 0.05), 0.2)).trigger()
    rep = C.Repeater(env)
    env.onFinish = lambda: setattr(rep, "IN", C.Osc(440))
    with pytest.raises(capi.UnsupportedCircuit, match="inside a tick"):
        circuit_blob(rep)
    osc = C.Osc(200)
    env = C.Shape("decay", 0.05).trigger()
    out = C.Multiply(osc, env)  # the oscillator ticks before the envelope
    env.onFinish = lambda: setattr(osc, "F", 300)
    with pytest.raises(capi.UnsupportedCircuit, match="already ticked"):
        circuit_blob(out)
    from dusp_b200 import patches as patches_
    assert "EPOCH" not in circuit_blob(patches_.Boop(440, 0.2))[0]


DEVICE_TRANSCENDENTAL = {"semi", "db2s", "midi2f", "pow", "gain", "pan", "filter"}  # libdevice pow / tan in the op


@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_exact_flag_follows_the_op_set(lib, name):
    """A case is held to bit-exactness iff no program of its timeline evaluates a transcendental on the device."""
    fac, _, exact = circuits.CASES[name]
    d.config.seed(12345)
    blob, _ = circuit_blob(fac())
    run = capi.Run([blob], [1])
    ops = {l.split()[1] for _, p in run.epochs(0) for l in p.disassemble().splitlines()[1:] if len(l.split()) > 1}
    run.close()
    assert exact == (not (ops & DEVICE_TRANSCENDENTAL)), sorted(ops & DEVICE_TRANSCENDENTAL)


def test_run_splits_a_multi_section_blob(lib):
    """dusp_run_create (csrc/run.cc) owns the run-time protocol: EPOCH sections become programs with their first
    chunk, circuits whose channel counts grow get their warm-up ticks as one-chunk epochs, and nothing needs a GPU
    until the first render."""
    blob, _ = circuit_blob(circuits.for_then())              # `(O440 for 0.5) then O220`: one graph edit at chunk 86
    run = capi.Run([blob], [3])
    ep = run.epochs(0)
    assert [k for k, _ in ep] == [0, 86] and run.num_voices == 3 and run.pos == 0
    assert ep[0][1].num_units < ep[1][1].num_units           # the edit added the second oscillator
    run.close()
    blob, _ = circuit_blob(circuits.scary_patch())           # 3 warm-up ticks before every unit sees both channels
    run = capi.Run([blob], [1])
    ep = run.epochs(0)
    assert [k for k, _ in ep] == list(range(len(ep))) and len(ep) >= 2
    assert [p.warmup_ticks for _, p in ep] == [len(ep) - 1] * len(ep)
    assert ep[0][1].num_channels <= ep[-1][1].num_channels == run.num_channels
    run.close()
    two = capi.Run([circuit_blob(circuits.at_later())[0], circuit_blob(circuits.then_number())[0]], [2, 5])
    assert len(two.epochs(0)) == 1 and len(two.epochs(1)) == 2 and two.num_voices == 7
    two.close()
    with pytest.raises(d.DuspError):
        capi.Run(["DUSPGRAPH 1\nEND\n"], [1])


def test_no_cpu_fallback(lib):
    """Without a CUDA device session creation must fail, not render on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    blob, _ = circuit_blob(C.Osc(440))
    p = capi.Program(blob)
    with pytest.raises(d.DuspError, match="no CUDA device|CUDA"):
        capi.Session([p], [1])
    with pytest.raises(d.DuspError):
        d.renderChannelData(C.Osc(440), 0.01)
    run = capi.Run([blob], [1])
    with pytest.raises(d.DuspError, match="no CUDA device|CUDA"):
        run.render_host(256)
    run.close()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "dusp_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cc", ".cu", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "import oracle" not in text and "from oracle" not in text, f


def _listing(prog):
    """[(op, unit label, [input refs], [output refs])] of a program's disassembly."""
    rows = []
    for l in prog.disassemble().splitlines()[1:]:
        t = l.split()
        if len(t) < 3:
            continue
        arrow = t.index("->")
        end = next(i for i, x in enumerate(t) if x.startswith("[p="))
        rows.append((t[1], t[2], t[3:arrow], t[arrow + 1:end]))
    return rows


def test_evaluation_order_of_feed_forward_circuits(lib, monkeypatch):
    """compile.cc re-sequences a program depth first where the tick order cannot be observed: the
    same instructions (one per unit and channel), every slot written before it is read, far fewer slots; circuits with
    CircleBuffers / feedback keep the reference's tick order."""
    from dusp_b200 import workloads as W
    for i in (0, 3, 7, 11, 19):
        blob = circuit_blob(W.w4_circuit(i))[0]
        new = capi.Program(blob)
        monkeypatch.setenv("DUSP_KEEP_TICK_ORDER", "1")
        old = capi.Program(blob)
        monkeypatch.delenv("DUSP_KEEP_TICK_ORDER")
        a, b = _listing(new), _listing(old)
        assert sorted((op, u) for op, u, _, _ in a) == sorted((op, u) for op, u, _, _ in b)  # a permutation of the same ops
        assert a[-1][0] == "out"
        live = set()
        for op, u, ins, outs in a:  # every slot read holds a value written earlier in the new order
            for r in ins:
                if r.startswith("s") and r[1:].isdigit():
                    assert r in live, (i, op, u, r)
            live.update(o for o in outs if o.startswith("s"))
        n_new = int(new.disassemble().split(" slots")[0].split()[-1])
        n_old = int(old.disassemble().split(" slots")[0].split()[-1])
        assert n_new <= n_old and n_new <= 6, (i, n_new, n_old)
    for fac in (W.w3_circuit, W.w3b_circuit):  # rings shared between units / a cycle edge: order kept
        blob = circuit_blob(fac())[0]
        new = capi.Program(blob)
        monkeypatch.setenv("DUSP_KEEP_TICK_ORDER", "1")
        old = capi.Program(blob)
        monkeypatch.delenv("DUSP_KEEP_TICK_ORDER")
        assert new.disassemble() == old.disassemble()
