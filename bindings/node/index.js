// index.js -- drop-in replacement for the reference's src/renderChannelData.js (joelle-o-world/dusp).
//
// NOT RUN IN THIS IMAGE (no node, no node_api.h; see INTEGRATION.md).  It keeps the reference's own Unit / Patch /
// Circuit / unDusp objects for graph construction -- that code is untouched -- and replaces only the tick loop
// (renderChannelData.js:29-45 -> Circuit.tick -> every unit's _tick).  What it does is what dusp_b200/render.py
// does with the Python mirror, statement for statement, and nothing else:
//   1. `outlet.unit.circuit || new Circuit(outlet.unit)` (renderChannelData.js:20): the reference's own scheduler
//      orders the units;
//   2. the circuit's event queue (Circuit.js:57-65, Event.js:18-30) is run ONCE, ahead of time, on the live objects,
//      and what every event did is recorded: numeric unit-state assignments -> `EV` records, graph edits (`then`,
//      events that assign inlets) -> a new `EPOCH` section holding the circuit after the edit, the assignments of a
//      (Sporadic)Retriggerer's `target.trigger()` -> `TG` records;
//   3. the text crosses the C ABI (include/dusp_b200.h): `dusp_run_create` owns everything that follows from those
//      records (segments, pokes, adopted sessions, warm-up programs -- csrc/run.cc), `dusp_run_render` fills the
//      channel data.
//
//   const dusp = require('dusp')                      // the reference, unmodified
//   dusp.renderChannelData = require('dusp-b200')     // this file
//   const cd = await dusp.renderChannelData(dusp.unDusp('(O440 for 0.5) then O220'), 10)
'use strict'
let addon_ = null
function addon() { return addon_ || (addon_ = require('./build/Release/dusp_b200_addon.node')) }  // on first render: the blob writer works without it
let refDir = 'dusp'
let Circuit = null
function reference(mod) { return require(refDir + '/src/' + mod) }
function getCircuit() { return Circuit || (Circuit = reference('Circuit')) }

const SR = 44100, CHUNK = 256
const EVENT_HORIZON_S = 600      // self-rescheduling events are expanded up to this render time (render.py)

// ---------------------------------------------------------------- the injected random generator
// Math.random inside Noise (Noise.js:10,23) and SporadicRetriggerer (SporadicRetrigger.js:13) is replaced on BOTH
// sides by a counter-based generator, draw k of stream `seed` = dusp_rand(seed, k) (include/dusp_b200.h): 32 bits,
// lowbias32(low32(k) + low32(seed)) ^ high32(seed), as a fraction of 2^32.  The splitmix below only hands out seeds.
const M64 = (1n << 64n) - 1n
function mix64(z) {
  z &= M64
  z = ((z ^ (z >> 30n)) * 0xBF58476D1CE4E5B9n) & M64
  z = ((z ^ (z >> 27n)) * 0x94D049BB133111EBn) & M64
  return z ^ (z >> 31n)
}
function tickRand(seed, k) {
  seed = BigInt(seed)
  let x = Number((BigInt(k) + (seed & 0xffffffffn)) & 0xffffffffn)
  x ^= x >>> 16; x = Math.imul(x, 0x7feb352d) >>> 0
  x ^= x >>> 15; x = Math.imul(x, 0x846ca68b) >>> 0
  x ^= x >>> 16
  return ((x ^ Number((seed >> 32n) & 0xffffffffn)) >>> 0) / 4294967296
}
let hostSeed = 12345n, noiseOrdinal = 0n
function seed(s) { hostSeed = BigInt(s); noiseOrdinal = 0n }                      // dusp_b200/config.py: seed()
function nextNoiseSeed() { return mix64((hostSeed ^ 0xD1B54A32D192ED03n) + (noiseOrdinal++) * 0x9E3779B97F4A7C15n) }
// Every stochastic unit of the circuit gets its own stream, in circuit.units order, the first time it is seen.
// A Noise's held value becomes draw 0 of its stream (what its constructor's Math.random() stands for); draws 1..
// happen in _tick.  A SporadicRetriggerer draws once per tick from draw 0 on.
function assignStreams(units) {
  for (const u of units) {
    const k = u.constructor.name
    if ((k === 'Noise' || k === 'SporadicRetriggerer') && u.dusp_b200_seed === undefined) {
      Object.defineProperty(u, 'dusp_b200_seed', { value: nextNoiseSeed(), enumerable: false, writable: true })
      Object.defineProperty(u, 'dusp_b200_draw', { value: k === 'Noise' ? 1 : 0, enumerable: false, writable: true })
      if (k === 'Noise') u.y = 2 * tickRand(u.dusp_b200_seed, 0) - 1
    }
  }
}
// For cross-checking: lets the UNTOUCHED reference render draw from the same streams.  Wraps the two _tick methods
// so that Math.random() inside them is draw k of the unit's stream.  (tools: bindings/node/ref_dump.js)
function installSeededRandom() {
  const real = Math.random
  let current = null
  Math.random = () => current ? tickRand(current.dusp_b200_seed, current.dusp_b200_draw++) : real()
  for (const mod of ['components/Noise', 'components/SporadicRetrigger']) {
    const proto = reference(mod).prototype
    if (proto._tick.dusp_b200_wrapped) continue
    const tick = proto._tick
    proto._tick = function (clock) {
      if (this.dusp_b200_seed === undefined) assignStreams([this])
      current = this
      try { return tick.call(this, clock) } finally { current = null }
    }
    proto._tick.dusp_b200_wrapped = true
  }
}

// ---------------------------------------------------------------- circuit -> "DUSPGRAPH 1" text
function num(x) {
  if (typeof x === 'boolean') return x ? '1' : '0'
  x = Number(x)
  if (Number.isNaN(x)) return 'nan'
  if (!Number.isFinite(x)) return x > 0 ? 'inf' : '-inf'
  return Number.isInteger(x) && Math.abs(x) < 1e15 ? String(x) : x.toPrecision(17)
}

// the state the device starts from; field for field what blob_props() returns in dusp_b200/components
function unitProps(u) {
  const P = {}, S = {}
  const k = u.constructor.name
  const frame = () => { P.windowSize = u.windowSize; P.tickInterval = u.tickInterval }
  switch (k) {
    case 'Osc': P.phase = u.phase; S.waveform = u._waveform !== undefined ? u._waveform : u.waveform; break
    case 'MultiChannelOsc': S.waveform = u._waveform !== undefined ? u._waveform : u.waveform; break
    case 'Ramp': P.duration = u.duration; P.y0 = u.y0; P.y1 = u.y1; P.t = u.t; P.playing = +!!u.playing; break
    case 'Shape':
      P.t = u.t; P.playing = +!!u.playing; P.finished = +!!u.finished; S.shape = u._shape !== undefined ? u._shape : u.shape
      for (const e of ['leftEdge', 'rightEdge']) (typeof u[e] === 'string' ? S : P)[e] = u[e]
      break
    case 'AHD': P.state = u.state; P.playing = +!!u.playing; P.t = u.t; break
    case 'Timer': P.t = u.t; break
    case 'Noise': {
      const s = BigInt(u.dusp_b200_seed)
      P.phase = u.phase; P.seed_lo = Number(s & 0xffffffffn); P.seed_hi = Number(s >> 32n); P.draw = u.dusp_b200_draw; P.y = u.y
      break
    }
    case 'SporadicRetriggerer': {
      const s = BigInt(u.dusp_b200_seed)
      P.seed_lo = Number(s & 0xffffffffn); P.seed_hi = Number(s >> 32n); P.draw = u.dusp_b200_draw
      break
    }
    case 'Retriggerer': P.t = u.t; break
    case 'Filter': S.kind = u._kind !== undefined ? u._kind : u.kind; break
    case 'Pan': P.compensationDB = u.compensationDB; break
    case 'FixedMultiply': P.sf = u.sf === undefined ? NaN : u.sf; break
    case 'SampleRateRedux': P.timeSinceLastUpdate = u.timeSinceLastUpdate; break
    case 'Delay': P.maxDelay = Math.trunc(u.maxDelay); break
    case 'ReadBackDelay': P.bufferLength = Math.trunc(u.bufferLength); P.tBuffer = u.tBuffer; break
    case 'FixedDelay': case 'CombFilter': case 'AllPass':
      P.delayTimeInSamples = Math.trunc(u.delayTimeInSamples); P.tBuffer = u.tBuffer; break
    case 'CircleBufferReader': P.t = u.t; P.postWipe = +!!u.postWipe; break
    case 'CircleBufferWriter': P.t = u.t; P.preWipe = +!!u.preWipe; break
    // frame-rate units (components/spectral)
    case 'Hopper': P.hopSize = u.hopSize; P.frameSize = u.frameSize; P.t = u.t; P.tickInterval = u.tickInterval; break
    case 'Windower': frame(); S.windowKind = u.windowKind; break
    case 'FFT': case 'IFFT': frame(); break
    case 'UnHopper': P.hopSize = u.hopSize; P.windowSize = u.windowSize; P.t = u.t; P.tickInterval = u.tickInterval; break
    case 'ReChunk':
      P.inputInterval = u.inputInterval; P.outputInterval = u.outputInterval; P.t = u.t
      P.bufferSize = u.bufferSize; P.tickInterval = u.tickInterval
      break
    case 'HardLowPass': case 'HardHighPass': case 'BinShift': case 'SpectralSum': frame(); break
    case 'SpectralGate': frame(); P.invert = +!!u.invert; break
    case 'Augment': {
      frame()
      const keys = Object.keys(u.incrementMapping)  // JavaScript's own-property order, as Augment._tick iterates
      if (keys.length > 8) throw 'Augment with more than 8 mapping entries has no device implementation'
      P.nmap = keys.length
      keys.forEach((key, i) => { P['ratio' + i] = Number(key); P['gain' + i] = Number(u.incrementMapping[key]) })
      break
    }
  }
  return { P, S }
}

function circuitToBlob(circuit, outlet) {
  const units = circuit.units
  assignStreams(units)
  const index = new Map(units.map((u, i) => [u, i]))
  const buffers = []
  for (const u of units) if (u.buffer && u.buffer.lengthInSamples !== undefined && !buffers.includes(u.buffer)) buffers.push(u.buffer)
  const L = ['DUSPGRAPH 1']
  for (const b of buffers) L.push(`CB ${b.numberOfChannels} ${b.lengthInSamples}`)
  for (const u of units) {
    L.push(`U ${u.constructor.name} ${String(u.label).replace(/\s/g, '_')}`)
    const { P, S } = unitProps(u)
    if (u.buffer && buffers.includes(u.buffer)) P.buffer = buffers.indexOf(u.buffer)
    for (const k in P) L.push(`P ${k} ${num(P[k])}`)
    for (const k in S) L.push(`S ${k} ${S[k]}`)
    for (const inlet of u.inletsOrdered) {
      const mono = inlet.exposeAsMono ? 1 : 0
      if (inlet.connected) {
        const src = inlet.outlet
        if (!index.has(src.unit)) throw 'inlet ' + inlet.label + ' is fed from outside the circuit'
        L.push(`I ${inlet.name} ${mono} ${inlet.noData ? 'N' : 'K'} ${index.get(src.unit)} ${src.unit.outletsOrdered.indexOf(src)}`)
      } else {
        const ch = inlet.signalChunk.channelData  // constant inlets hold their value in every sample (Inlet.js:76-93)
        L.push(`I ${inlet.name} ${mono} C ${ch.length} ${Array.from(ch, c => num(c[0])).join(' ')}`)
      }
    }
    for (const o of u.outletsOrdered)
      L.push(`O ${o.name} ${o.exposeAsMono ? 1 : 0} ${o.signalChunk.channelData.length} ${o.connections.filter(i => index.has(i.unit)).map(i => index.get(i.unit)).join(' ')}`.trim())
  }
  L.push('ORDER ' + units.map((_, i) => i).join(' '))
  L.push(`ROOT ${index.get(outlet.unit)} ${outlet.unit.outletsOrdered.indexOf(outlet)}`)
  L.push('END')
  return L.join('\n') + '\n'
}

// ---------------------------------------------------------------- recording what an event does
// host-side bookkeeping an event may assign: finish flags and what Circuit.add / computeProcessIndex write while an
// edit adds units (Circuit.js:67-107, Unit.js:171-209)
const IGNORED = new Set(['finished', 'onFinish', 'processIndex', 'circuit', 'events', 'promises', 'nChains', 'dusp_b200_draw', 'dusp_b200_seed'])

// Runs f() with every own data property of every unit turned into an accessor that logs assignments:
// [[unit, property, value]] in program order (a later assignment of the same property wins when replayed).
function recordAssignments(units, f) {
  const rec = [], undo = []
  for (const u of units) {
    for (const k of Object.keys(u)) {
      const d = Object.getOwnPropertyDescriptor(u, k)
      if (!d || !d.configurable || !('value' in d) || typeof d.value === 'function') continue  // (objects too: an event that swaps a table or a buffer is refused, not ignored)
      let shadow = d.value
      Object.defineProperty(u, k, { configurable: true, enumerable: d.enumerable, get: () => shadow, set: v => { rec.push([u, k, v]); shadow = v } })
      undo.push(() => Object.defineProperty(u, k, { configurable: true, enumerable: d.enumerable, writable: true, value: shadow }))
    }
    if (u.constructor.name === 'MultiChannelOsc') {  // resetPhase assigns array ELEMENTS (MultiChannelOsc.js:52-55): every channel that exists by then
      u.resetPhase = function () { rec.push([u, 'phase[*]', 0]); return Object.getPrototypeOf(u).resetPhase.call(u) }
      undo.push(() => { delete u.resetPhase })
    }
  }
  try { f() } finally { for (const g of undo.reverse()) g() }
  return rec
}

function numericAssignments(rec, what) {
  const out = []
  for (const [unit, name, val] of rec) {
    if (IGNORED.has(name) || name.startsWith('_') || name.startsWith('@')) continue
    let v = val
    if (typeof v === 'boolean') v = v ? 1 : 0
    else if (typeof v !== 'number') throw `${what} assigns ${unit.label}.${name}: only numeric unit state and inlets can change during an accelerated render`
    out.push([unit, name, v])
  }
  return out
}

// what a graph edit changes: the unit list and every inlet's source / constant
function structure(circuit) {
  return JSON.stringify(circuit.units.map(u => [u.label, u.inletsOrdered.map(inlet => inlet.connected
    ? [inlet.name, inlet.outlet.unit.label, inlet.outlet.name]
    : [inlet.name, Array.from(inlet.signalChunk.channelData, c => num(c[0]))])]))
}

// Later events (render.py: collect_timeline).  Returns [[startChunk, graphText, units, [[chunk, unitIndex, property, value]]]].
function collectTimeline(circuit, outlet, horizonSeconds) {
  const chunk = circuit.gcdTickInterval, horizon = horizonSeconds * SR
  const epochs = [[0, circuitToBlob(circuit, outlet), circuit.units.slice()]]
  const pending = []
  let sig = structure(circuit), guard = 0, kNow = 0
  while (circuit.events.length && circuit.events[0].t < horizon) {
    const ev = circuit.events.shift()
    // (an event queued in the past by a unit that joined the circuit late runs in the same runEvents pass)
    const k = kNow = Math.max(kNow, Math.floor(ev.t / chunk))
    const before = new Set(circuit.units)
    let follow = null
    const rec = recordAssignments(circuit.units.slice(), () => { follow = ev.run() }).filter(r => before.has(r[0]))
    for (const a of numericAssignments(rec, 'event at t=' + ev.t)) pending.push([k, ...a])
    const now = structure(circuit)
    if (now !== sig) {
      sig = now
      const snap = [k, circuitToBlob(circuit, outlet), circuit.units.slice()]
      if (k === epochs[epochs.length - 1][0] && epochs.length > 1) epochs[epochs.length - 1] = snap
      else if (k === 0) throw 'graph edit before the first chunk'
      else epochs.push(snap)
    }
    if (follow) circuit.addEvent(follow)
    if (++guard > 200000) throw 'more than 200000 events inside the event horizon'
  }
  if (epochs.length > 1)
    for (const [, , units] of epochs)
      if (new Set(units.map(u => u.label)).size !== units.length) throw 'graph edits need unique unit labels (units are matched by label)'
  return epochs.map(([k0, text, units], e) => {
    const k1 = e + 1 < epochs.length ? epochs[e + 1][0] : null
    const index = new Map(units.map((u, i) => [u, i]))
    const evs = []
    for (const [k, unit, name, val] of pending) {
      if (k < k0 || (k1 !== null && k >= k1)) continue
      if (!index.has(unit)) throw 'event touches a unit outside the rendered circuit: ' + unit.label
      evs.push([k, index.get(unit), name, val])
    }
    return [k0, text, units, evs]
  })
}

// (Sporadic)Retriggerer units (Retriggerer.js:17-18): `target.trigger()` is run ONCE with its assignments recorded;
// the device and the oracle replay them whenever the retriggerer fires.  [[retriggererIndex, unitIndex, property, value]]
function collectTriggers(units) {
  const index = new Map(units.map((u, i) => [u, i]))
  const out = []
  units.forEach((u, i) => {
    const k = u.constructor.name
    if (k !== 'Retriggerer' && k !== 'SporadicRetriggerer') return
    const target = u.target
    if (!target || typeof target.trigger !== 'function') return
    const rec = recordAssignments(units, () => target.trigger())
    for (const [unit, name, val] of numericAssignments(rec, 'trigger() of ' + target.label)) {
      if (!index.has(unit)) throw `trigger() of ${target.label} touches a unit outside the rendered circuit: ${unit.label}`
      out.push([i, index.get(unit), name, val])
    }
  })
  return out
}

// A Shape calls finish() from inside its _tick when its envelope runs out (Shape/index.js:40-42): an onFinish that
// edits the graph or assigns unit state there acts in the middle of Circuit.tick.  (render.py: _reject_mid_tick_finish)
function rejectMidTickFinish(circuit, units) {
  for (const u of units) {
    if (u.constructor.name !== 'Shape' || typeof u.onFinish !== 'function') continue
    const sig = structure(circuit)
    const rec = recordAssignments(circuit.units.slice(), () => u.onFinish())
    if (structure(circuit) !== sig || numericAssignments(rec, 'onFinish of ' + u.label).length)
      throw `onFinish of ${u.label} would act from inside a tick when the envelope runs out (Shape/index.js:40-42): outside the accelerated render path`
  }
}

// (blob, circuit) for a rendered outlet -- render.py: circuit_blob
function circuitBlob(outlet, horizonSeconds = EVENT_HORIZON_S) {
  if (!outlet) throw 'renderAudioBuffer expects an outlet'                       // renderChannelData.js:12-17
  if (outlet.isUnit || outlet.isPatch) outlet = outlet.defaultOutlet
  if (!outlet || !outlet.isOutlet) throw 'renderAudioBuffer expects an outlet'
  const circuit = outlet.unit.circuit || new (getCircuit())(outlet.unit)       // renderChannelData.js:20
  if (circuit.dusp_b200_rendered) throw 'this circuit has already been rendered: renders are one-shot'
  circuit.runEvents(circuit.clock + circuit.gcdTickInterval)                    // events before the first tick fold into state
  const epochs = collectTimeline(circuit, outlet, horizonSeconds)
  const parts = []
  epochs.forEach(([k0, text, units, evs], e) => {
    const lines = evs.map(ev => `EV ${ev[0]} ${ev[1]} ${ev[2]} ${num(ev[3])}`)
      .concat(collectTriggers(units).map(t => `TG ${t[0]} ${t[1]} ${t[2]} ${num(t[3])}`))
    if (e + 1 < epochs.length) lines.push('KEEP')  // an edit follows: the program keeps every outlet's last chunk for its successor
    if (lines.length) text = text.slice(0, -'END\n'.length) + lines.join('\n') + '\nEND\n'
    parts.push((k0 ? `EPOCH ${k0}\n` : '') + text)
  })
  const all = new Set()
  for (const [, , units] of epochs) for (const u of units) all.add(u)
  rejectMidTickFinish(circuit, all)
  return { blob: parts.join(''), circuit, outlet }
}

// ---------------------------------------------------------------- the sink
// renderChannelData.js:5-49.  Same argument checks and thrown strings; resolves to Array<TypedArray> & {sampleRate}.
async function renderChannelData(outlet, duration = 1, { TypedArray = Float32Array, device = 0 } = {}) {
  const { blob, circuit, outlet: out } = circuitBlob(outlet)
  const length = duration * out.sampleRate                                       // renderChannelData.js:24
  const n = Number.isFinite(length) ? Math.max(0, Math.trunc(length)) : 0
  const { data, channels } = await addon().renderRun([blob], [1], [], [], n, 0, device)
  circuit.dusp_b200_rendered = true
  const channelData = []
  for (let c = 0; c < channels; c++) channelData.push(TypedArray.from(data.subarray(c * n, (c + 1) * n)))
  channelData.sampleRate = out.sampleRate
  return channelData
}

// Many voices of ONE circuit shape in one launch (dusp_b200/render.py: Batch).  `overrides` maps parameter names
// ("u<unit>.<inlet>[<ch>]" / "u<unit>.<prop>") to one value per voice.  Resolves to a Float32Array laid out
// [voice][channel][n] (outMode 0) or [2][n] (outMode 1, the stereo mixdown) plus the channel count.
async function renderBatch(outlet, nVoices, duration = 1, { overrides = {}, outMode = 0, device = 0 } = {}) {
  const { blob, circuit, outlet: out } = circuitBlob(outlet)
  const names = Object.keys(overrides)
  const values = new Float64Array(nVoices * names.length)
  names.forEach((nm, k) => {
    if (overrides[nm].length !== nVoices) throw 'override arrays must hold one value per voice'
    for (let v = 0; v < nVoices; v++) values[v * names.length + k] = overrides[nm][v]
  })
  const n = Math.max(0, Math.trunc(duration * out.sampleRate))
  const res = await addon().renderRun([blob], [nVoices], [names], [values], n, outMode, device)
  circuit.dusp_b200_rendered = true
  return res
}

module.exports = renderChannelData
Object.assign(module.exports, {
  renderChannelData, renderBatch, circuitBlob, circuitToBlob, collectTimeline, collectTriggers, assignStreams,
  installSeededRandom, seed, tickRand,
  useReference(dir) { refDir = dir; Circuit = null },
})
