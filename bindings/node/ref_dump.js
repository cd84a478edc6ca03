// ref_dump.js -- runs the UNTOUCHED reference (joelle-o-world/dusp) on a Node-equipped host and writes raw float32
// renders of EVERY parity case (bindings/node/cases.json, generated from tests/circuits.py by
// tests/golden/make_cases_json.py), so that the C oracle (oracle/dusp_oracle.c) can be pinned against the reference
// itself instead of against closed forms only (DESIGN.md section 2, "parity unpinned").
//
// NOT RUN IN THIS IMAGE: node is absent here and on the GPU box.  One command on a host that has it:
//   git clone https://github.com/joelle-o-world/dusp && (cd dusp && npm i minimist@1.2.0 promise@8.0.2 compute-gcd@1.2.0 fft.js@4.0.3)
//   node bindings/node/ref_dump.js ./dusp tests/golden/ref_out && python tests/golden/import_ref.py tests/golden/ref_out
// writes <out>/<case>.f32 ([channel][n] little-endian) + <out>/index.json; import_ref.py turns that into
// tests/golden/golden_ref.npz, which tests/test_golden.py::test_oracle_matches_reference_renders then checks.
//
// A case is either Dusp text (parsed by the reference's own unDusp) or a "DUSPGRAPH 1" blob, from which the circuit
// is rebuilt here out of the reference's own components: one `new components.<Kind>(...)` per U record, properties,
// inlet constants / connections by name, chain edges.  Seeded noise: Math.random inside Noise / SporadicRetriggerer
// ticks is the injected counter-based generator (dusp_rand, include/dusp_b200.h), every unit with its own stream.
'use strict'
const fs = require('fs'), path = require('path')
const [refDirArg, outDir] = process.argv.slice(2)
if (!refDirArg || !outDir) { console.error('usage: node ref_dump.js /path/to/dusp out_dir [case ...]'); process.exit(2) }
const refDir = path.resolve(refDirArg)
const only = process.argv.slice(4)
const dusp = require(path.join(refDir, 'src'))
const Circuit = require(path.join(refDir, 'src/Circuit'))
const CircleBuffer = require(path.join(refDir, 'src/CircleBuffer'))
const C = dusp.components
const SR = 44100

// ---- the injected generator (host seed -> per-unit stream seeds -> draws), as dusp_b200/config.py
const M64 = (1n << 64n) - 1n
function mix64(z) {
  z &= M64
  z = ((z ^ (z >> 30n)) * 0xBF58476D1CE4E5B9n) & M64
  z = ((z ^ (z >> 27n)) * 0x94D049BB133111EBn) & M64
  return z ^ (z >> 31n)
}
function tickRand(seed, k) {  // dusp_rand: lowbias32(low32(k) + low32(seed)) ^ high32(seed), as a fraction of 2^32
  let x = Number((BigInt(k) + (seed & 0xffffffffn)) & 0xffffffffn)
  x ^= x >>> 16; x = Math.imul(x, 0x7feb352d) >>> 0
  x ^= x >>> 15; x = Math.imul(x, 0x846ca68b) >>> 0
  x ^= x >>> 16
  x = (x ^ Number(seed >> 32n)) >>> 0
  return x / 4294967296
}
let hostSeed = 12345n, noiseOrdinal = 0n
function nextNoiseSeed() { return mix64((hostSeed ^ 0xD1B54A32D192ED03n) + (noiseOrdinal++) * 0x9E3779B97F4A7C15n) }
let current = null
const realRandom = Math.random
Math.random = () => current ? tickRand(current.seed, current.k++) : realRandom()
function giveStream(u, seed, draw) {  // draws inside this unit's _tick come from its own stream
  const st = { seed: BigInt(seed), k: draw }
  const tick = u._tick
  u._tick = function (clock) { current = st; try { return tick.call(this, clock) } finally { current = null } }
  return st
}
function seedStochasticUnits(units) {  // Dusp-text cases: streams in circuit.units order, as the Python mirror hands them out
  for (const u of units) {
    const k = u.constructor.name
    if (k === 'Noise') { const s = nextNoiseSeed(); giveStream(u, s, 1); u.y = 2 * tickRand(s, 0) - 1 }
    else if (k === 'SporadicRetriggerer') giveStream(u, nextNoiseSeed(), 0)
  }
}
// (every Dusp-text case holds at most one stochastic unit, so construction order and circuit order agree)

// ---- "DUSPGRAPH 1" blob -> the reference's own objects
function parseBlob(text) {
  const g = { cbs: [], units: [], order: null, root: null, evs: [] }
  let u = null
  for (const line of text.split('\n')) {
    const t = line.trim().split(/\s+/)
    if (!t[0] || t[0] === 'DUSPGRAPH' || t[0] === 'KEEP' || t[0] === 'WARMUP') continue
    if (t[0] === 'END') break
    if (t[0] === 'CB') g.cbs.push({ nch: +t[1], len: +t[2] })
    else if (t[0] === 'U') { u = { kind: t[1], label: t[2], P: {}, S: {}, ins: [], outs: [] }; g.units.push(u) }
    else if (t[0] === 'P') u.P[t[1]] = t[2] === 'nan' ? NaN : t[2] === 'inf' ? Infinity : t[2] === '-inf' ? -Infinity : Number(t[2])
    else if (t[0] === 'S') u.S[t[1]] = t[2]
    else if (t[0] === 'I') {
      if (t[3] === 'C') u.ins.push({ name: t[1], consts: t.slice(5, 5 + +t[4]).map(x => x === 'nan' ? NaN : Number(x)) })
      else u.ins.push({ name: t[1], chain: t[3] === 'N', src: +t[4], outlet: +t[5] })
    } else if (t[0] === 'O') u.outs.push({ name: t[1], nch: +t[3] })
    else if (t[0] === 'ORDER') g.order = t.slice(1).map(Number)
    else if (t[0] === 'ROOT') g.root = { unit: +t[1], outlet: +t[2] }
    else if (t[0] === 'EV') g.evs.push({ chunk: +t[1], unit: +t[2], prop: t[3], value: Number(t[4]) })
    else if (t[0] === 'TG') throw 'Retriggerer targets cannot be rebuilt from a blob'
    else throw 'unknown blob record ' + t[0]
  }
  return g
}

function construct(u, buffers) {  // constructor arguments that allocate or size something; everything else is set afterwards
  const P = u.P, K = C[u.kind] || dusp.components[u.kind]
  if (!K) throw 'no reference component ' + u.kind
  switch (u.kind) {
    case 'Osc': case 'MultiChannelOsc': return new K(undefined, u.S.waveform)
    case 'Shape': return new K(u.S.shape)
    case 'Filter': return new K(undefined, undefined, u.S.kind)
    case 'Delay': return new K(undefined, undefined, P.maxDelay)
    case 'ReadBackDelay': return new K(undefined, undefined, P.bufferLength)
    case 'FixedDelay': case 'CombFilter': case 'AllPass': {
      const x = new K((P.delayTimeInSamples - 0.25) / SR)   // setSeconds rounds to samples (FixedDelay.js:22-28)
      if (x.delayTimeInSamples !== P.delayTimeInSamples) throw u.kind + ': ring of ' + x.delayTimeInSamples + ' instead of ' + P.delayTimeInSamples
      return x
    }
    case 'CircleBufferReader': case 'CircleBufferWriter': return new K(buffers[P.buffer])
    case 'FixedMultiply': return new K(P.sf)
    case 'Hopper': return new K(P.hopSize, P.frameSize)
    case 'Windower': return new K(P.windowSize, u.S.windowKind, P.tickInterval)
    case 'FFT': case 'IFFT': return new K(P.windowSize, P.tickInterval)
    case 'UnHopper': return new K(P.hopSize, P.windowSize)
    case 'ReChunk': return new K(P.inputInterval, P.outputInterval)
    case 'Augment': {
      const m = {}
      for (let i = 0; i < P.nmap; i++) m[String(P['ratio' + i])] = P['gain' + i]
      return new K(m)
    }
    default: return new K()
  }
}

const STATE = ['phase', 't', 'playing', 'y0', 'y1', 'duration', 'state', 'finished', 'tBuffer', 'compensationDB', 'timeSinceLastUpdate',
  'preWipe', 'postWipe', 'invert', 'leftEdge', 'rightEdge']
const BOOLS = new Set(['playing', 'finished', 'preWipe', 'postWipe', 'invert'])

function fromBlob(text) {
  const g = parseBlob(text)
  const buffers = g.cbs.map(b => {
    const cb = new CircleBuffer(b.nch, (b.len - 0.5) / SR)
    if (cb.lengthInSamples !== b.len) throw 'CircleBuffer of ' + cb.lengthInSamples + ' samples instead of ' + b.len
    return cb
  })
  const units = g.units.map(u => construct(u, buffers))
  g.units.forEach((u, i) => {
    const x = units[i]
    x.label = u.label
    for (const k of STATE) if (k in u.P) x[k] = BOOLS.has(k) ? !!u.P[k] : u.P[k]
    for (const k of ['leftEdge', 'rightEdge']) if (k in u.S) x[k] = u.S[k]
    if (u.kind === 'Noise') { giveStream(x, BigInt(u.P.seed_lo) | (BigInt(u.P.seed_hi) << 32n), u.P.draw); x.phase = u.P.phase; x.y = u.P.y }
    if (u.kind === 'SporadicRetriggerer') giveStream(x, BigInt(u.P.seed_lo) | (BigInt(u.P.seed_hi) << 32n), u.P.draw)
    u.outs.forEach((o, k) => {   // an outlet whose channel count was set by hand (multi-channel Noise ...)
      const cd = x.outletsOrdered[k].signalChunk.channelData
      while (cd.length < o.nch) cd.push(new Float32Array(cd[0].length))
    })
  })
  g.units.forEach((u, i) => {   // constants first, then connections in circuit order (outlet.connections order feeds the scheduler)
    for (const inl of u.ins) if (inl.consts) units[i][inl.name.toUpperCase()] = inl.consts.length === 1 ? inl.consts[0] : inl.consts
  })
  g.units.forEach((u, i) => {
    for (const inl of u.ins) {
      if (inl.consts) continue
      if (inl.chain) units[i].chainAfter(units[inl.src])
      else units[i][inl.name.toUpperCase()] = units[inl.src].outletsOrdered[inl.outlet]
    }
  })
  const rootUnit = units[g.root.unit]
  const circuit = new Circuit(rootUnit)
  // The reference's tick order depends on construction history (Unit.js:171-209 is not a pure function of the graph).
  // The blob's ORDER is the order the original construction produced: impose it if the rebuild came out differently.
  const want = g.order.map(i => units[i])
  if (want.length === circuit.units.length && want.some((x, i) => x !== circuit.units[i])) {
    console.warn('   (tick order imposed from the blob)')
    circuit.units = want
  }
  for (const ev of g.evs) {   // the assignments events make, at the chunk the host mirror found them at
    const x = units[ev.unit]
    x.schedule(ev.chunk * 256 / SR, function () {
      if (ev.prop === 'phase[*]') { for (const i in this.phase) this.phase[i] = ev.value } else this[ev.prop] = BOOLS.has(ev.prop) ? !!ev.value : ev.value
    })
  }
  return rootUnit.outletsOrdered[g.root.outlet]
}

;(async () => {
  fs.mkdirSync(outDir, { recursive: true })
  const cases = JSON.parse(fs.readFileSync(path.join(__dirname, 'cases.json'), 'utf8'))
  const index = {}, failed = {}
  for (const name of Object.keys(cases).sort()) {
    if (only.length && !only.includes(name)) continue
    const c = cases[name]
    if (c.skip) { index[name] = { skipped: c.skip }; continue }
    try {
      hostSeed = BigInt(c.seed); noiseOrdinal = 0n
      let outlet
      if (c.dusp !== undefined) {
        const x = dusp.unDusp(c.dusp)
        outlet = x.isOutlet ? x : x.defaultOutlet
        const circuit = outlet.unit.circuit || new Circuit(outlet.unit)
        seedStochasticUnits(circuit.units)
      } else outlet = fromBlob(c.blob)
      const cd = await dusp.renderChannelData(outlet, c.seconds)
      const n = cd.length ? cd[0].length : 0
      const flat = new Float32Array(cd.length * n)
      cd.forEach((ch, i) => flat.set(ch, i * n))
      fs.writeFileSync(path.join(outDir, name + '.f32'), Buffer.from(flat.buffer))
      index[name] = { channels: cd.length, samples: n, seconds: c.seconds, exact: c.exact, from: c.dusp !== undefined ? 'dusp' : 'blob' }
      console.log('ok  ', name, cd.length + ' x ' + n)
    } catch (e) {
      failed[name] = String(e && e.stack || e)
      console.log('FAIL', name, String(e))
    }
  }
  fs.writeFileSync(path.join(outDir, 'index.json'), JSON.stringify({ cases: index, failed }, null, 1))
  console.log(Object.keys(index).length + ' cases written, ' + Object.keys(failed).length + ' failed')
})()
