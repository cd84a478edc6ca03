// index.js -- drop-in replacement for the reference's src/renderChannelData.js.
//
// NOT RUN IN THIS IMAGE (no node).  It keeps the reference's own Unit / Patch /
// Circuit / unDusp objects for graph construction -- that code is untouched --
// and replaces only the tick loop: the circuit is serialised to the
// "DUSPGRAPH 1" text the C ABI takes (same format as dusp_b200/graphblob.py)
// and rendered by libdusp_b200 through the N-API addon.
//
//   const dusp = require('dusp')                      // the reference, unmodified
//   dusp.renderChannelData = require('dusp-b200')     // this file
//   const cd = await dusp.renderChannelData(dusp.unDusp('[Osc 440]'), 10)
const addon = require('./build/Release/dusp_b200_addon.node')
const Circuit = require('dusp/src/Circuit')

const KINDS = {  // reference constructor.name -> DUSPGRAPH kind
  SporadicRetriggerer: null, Retriggerer: null, Monitor: null,  // run-time graph edits: unsupported
}

function num(x) { return Number(x).toPrecision(17) }

function unitProps(u) {  // the state the device starts from; mirrors blob_props() in dusp_b200/components
  const P = {}, S = {}
  const k = u.constructor.name
  if (k === 'Osc') { P.phase = u.phase; S.waveform = u.waveform }
  if (k === 'MultiChannelOsc') { S.waveform = u.waveform }
  if (k === 'Ramp') { P.duration = u.duration; P.y0 = u.y0; P.y1 = u.y1; P.t = u.t; P.playing = +u.playing }
  if (k === 'Shape') {
    P.t = u.t; P.playing = +u.playing; P.finished = +!!u.finished; S.shape = u.shape
    for (const e of ['leftEdge', 'rightEdge']) (typeof u[e] === 'string' ? S : P)[e] = u[e]
  }
  if (k === 'AHD') { P.state = u.state; P.playing = +u.playing; P.t = u.t }
  if (k === 'Timer') { P.t = u.t }
  if (k === 'Noise') {  // the injected generator: dusp_b200_seed is attached by the shim's Math.random wrapper
    const s = BigInt(u.dusp_b200_seed || 0)
    P.phase = u.phase; P.y = u.y; P.draw = 1
    P.seed_lo = Number(s & 0xffffffffn); P.seed_hi = Number(s >> 32n)
  }
  if (k === 'Filter') { S.kind = u.kind }
  if (k === 'Pan') { P.compensationDB = u.compensationDB }
  if (k === 'FixedMultiply') { P.sf = u.sf }
  if (k === 'Delay') { P.maxDelay = u.maxDelay }
  if (k === 'ReadBackDelay') { P.bufferLength = u.bufferLength; P.tBuffer = u.tBuffer }
  if (k === 'FixedDelay' || k === 'CombFilter' || k === 'AllPass') { P.delayTimeInSamples = u.delayTimeInSamples; P.tBuffer = u.tBuffer }
  if (k === 'CircleBufferReader') { P.t = u.t; P.postWipe = +!!u.postWipe }
  if (k === 'CircleBufferWriter') { P.t = u.t; P.preWipe = +!!u.preWipe }
  return { P, S }
}

function circuitToBlob(circuit, outlet) {
  const units = circuit.units
  const index = new Map(units.map((u, i) => [u, i]))
  const buffers = []
  for (const u of units) if (u.buffer && !buffers.includes(u.buffer)) buffers.push(u.buffer)
  const L = ['DUSPGRAPH 1']
  for (const b of buffers) L.push(`CB ${b.numberOfChannels} ${b.lengthInSamples}`)
  for (const u of units) {
    const kind = u.constructor.name
    if (KINDS[kind] === null) throw 'unit kind has no device implementation: ' + kind
    L.push(`U ${kind} ${String(u.label).replace(/\s/g, '_')}`)
    const { P, S } = unitProps(u)
    if (u.buffer) P.buffer = buffers.indexOf(u.buffer)
    for (const k in P) L.push(`P ${k} ${num(P[k])}`)
    for (const k in S) L.push(`S ${k} ${S[k]}`)
    for (const inlet of u.inletsOrdered) {
      const mono = inlet.exposeAsMono ? 1 : 0
      if (inlet.connected) {
        const src = inlet.outlet
        L.push(`I ${inlet.name} ${mono} ${inlet.noData ? 'N' : 'K'} ${index.get(src.unit)} ${src.unit.outletsOrdered.indexOf(src)}`)
      } else {
        const ch = inlet.signalChunk.channelData  // constant inlets hold their value in every sample (Inlet.js:76-93)
        L.push(`I ${inlet.name} ${mono} C ${ch.length} ${ch.map(c => num(c[0])).join(' ')}`)
      }
    }
    for (const o of u.outletsOrdered)
      L.push(`O ${o.name} ${o.exposeAsMono ? 1 : 0} ${o.signalChunk.channelData.length} ${o.connections.map(i => index.get(i.unit)).join(' ')}`.trim())
  }
  L.push('ORDER ' + units.map((_, i) => i).join(' '))
  L.push(`ROOT ${index.get(outlet.unit)} ${outlet.unit.outletsOrdered.indexOf(outlet)}`)
  L.push('END')
  return L.join('\n') + '\n'
}

async function renderChannelData(outlet, duration = 1, { TypedArray = Float32Array } = {}) {
  if (!outlet) throw 'renderAudioBuffer expects an outlet'
  if (outlet.isUnit || outlet.isPatch) outlet = outlet.defaultOutlet
  if (!outlet.isOutlet) throw 'renderAudioBuffer expects an outlet'
  const circuit = outlet.unit.circuit || new Circuit(outlet.unit)
  circuit.runEvents(circuit.clock + circuit.gcdTickInterval)  // events before the first tick fold into state
  if (circuit.events.length) throw 'events after the first chunk are outside the accelerated render path'
  const n = Math.max(0, Math.trunc(duration * outlet.sampleRate))
  const flat = await addon.renderBatch(circuitToBlob(circuit, outlet), 1, [], new Float64Array(0), n, 0)
  const nch = n ? flat.length / n : 0
  const channelData = []
  for (let c = 0; c < nch; c++) channelData.push(TypedArray.from(flat.subarray(c * n, (c + 1) * n)))
  channelData.sampleRate = outlet.sampleRate
  return channelData
}

module.exports = renderChannelData
module.exports.circuitToBlob = circuitToBlob
module.exports.renderBatch = addon.renderBatch
