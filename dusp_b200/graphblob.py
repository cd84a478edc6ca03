"""Circuit -> "DUSPGRAPH 1" text blob: the neutral graph description that
crosses the C-ABI (``dusp_program_create`` in ``include/dusp_b200.h``).  A
Node.js shim would emit the same text from the reference's own objects
(INTEGRATION.md); the parity oracle reads it too.

Format (one record per line, space separated, doubles as ``%.17g``):

  DUSPGRAPH 1
  CB <nChannels> <lengthInSamples>            shared CircleBuffer (index = order of appearance)
  U <Kind> <label>                            unit (index = order of appearance)
  P <name> <number>                           numeric property / initial state of the last U
  S <name> <string>                           string property of the last U
  I <name> <mono> C <n> <v0> .. <v(n-1)>      unconnected inlet: per-channel chunk constants
  I <name> <mono> K <srcUnit> <srcOutlet>     connected inlet (Inlet.connect)
  I <name> <mono> N <srcUnit> <srcOutlet>     order-only "chain" connection (Unit.chainAfter)
  O <name> <mono> <nChannels> [<unit> ...]    outlet, its initial channel count, and the units of
                                              outlet.connections in connect order (scheduler input)
  ORDER <u0> <u1> ...                         circuit.units (tick order)
  ROOT <unit> <outlet>                        the rendered outlet
  EV <chunk> <unit> <prop> <number>           an event's assignment, applied before the tick of chunk <chunk>
  TG <unit> <target> <prop> <number>          one assignment of target.trigger() for (Sporadic)Retriggerer <unit>
  KEEP                                        a graph edit follows: keep every outlet's last chunk (device hint)
  END
  EPOCH <chunk>                               then another DUSPGRAPH section: the circuit after an event edited the
                                              graph, valid from that chunk on (units matched by kind and label)
"""
from .graph import DuspError


def _num(x):
    if isinstance(x, bool):
        return "1" if x else "0"
    if isinstance(x, int):
        return repr(x)
    return "%.17g" % float(x)


def circuit_to_blob(circuit, root_outlet):
    units = list(circuit.units)
    # Index units in a construction-independent but deterministic way: the
    # circuit's own unit list order.  ORDER then is simply 0..n-1, but it is
    # written explicitly because the library never re-derives it.
    index = {id(u): i for i, u in enumerate(units)}
    # outlet.connections order matters to Unit.outputUnits (Unit.js:133-143);
    # the blob lists units in circuit order and inlets in declaration order, and
    # readers that re-run the scheduler rebuild `connections` in that order.
    buffers = []
    lines = ["DUSPGRAPH 1"]
    for u in units:
        b = getattr(u, "buffer", None)
        if b is not None and all(b is not x for x in buffers):
            buffers.append(b)
    for b in buffers:
        lines.append("CB %d %d" % (b.numberOfChannels, b.lengthInSamples))
    for u in units:
        if u.KIND is None:
            raise DuspError("unit kind has no device implementation: " + type(u).__name__)
        lines.append("U %s %s" % (u.KIND, u.label.replace(" ", "_")))
        nums, strs = u.blob_props()
        b = getattr(u, "buffer", None)
        if b is not None:
            nums = dict(nums)
            nums["buffer"] = [i for i, x in enumerate(buffers) if x is b][0]
        for k, v in nums.items():
            lines.append("P %s %s" % (k, _num(v)))
        for k, v in strs.items():
            lines.append("S %s %s" % (k, v))
        for inlet in u.inletsOrdered:
            mono = 1 if inlet.exposeAsMono else 0
            if inlet.connected:
                src = inlet.outlet
                if id(src.unit) not in index:
                    raise DuspError("inlet " + inlet.label + " is fed from outside the circuit")
                lines.append("I %s %d %s %d %d" % (inlet.name, mono, "N" if inlet.noData else "K",
                                                   index[id(src.unit)], src.unit.outletsOrdered.index(src)))
            else:
                lines.append("I %s %d C %d %s" % (inlet.name, mono, len(inlet.chunk),
                                                  " ".join(_num(v) for v in inlet.chunk)))
        for outlet in u.outletsOrdered:
            conns = " ".join(str(index[id(i.unit)]) for i in outlet.connections if id(i.unit) in index)
            lines.append(("O %s %d %d %s" % (outlet.name, 1 if outlet.exposeAsMono else 0, len(outlet.chunk),
                                              conns)).rstrip())
    lines.append("ORDER " + " ".join(str(i) for i in range(len(units))))
    lines.append("ROOT %d %d" % (index[id(root_outlet.unit)], root_outlet.unit.outletsOrdered.index(root_outlet)))
    lines.append("END")
    return "\n".join(lines) + "\n"


def strip_order(blob):
    """Blob without its ORDER line (lets the oracle run its own restatement of
    ``new Circuit(root)`` -- used to cross-check the host scheduler)."""
    return "".join(l + "\n" for l in blob.splitlines() if not l.startswith("ORDER"))
