// ref_dump.js -- runs the UNTOUCHED reference (joelle-o-world/dusp) on a Node-equipped host and writes raw
// float32 renders of the parity cases, so the C oracle (oracle/dusp_oracle.c) can be pinned against the
// reference itself instead of against closed forms only (DESIGN.md section 2, "parity unpinned").
//
// NOT RUN IN THIS IMAGE: node is absent here and on the GPU box.
//   npm i minimist@1.2.0 promise@8.0.2 compute-gcd@1.2.0      (the reference's un-vendored deps on this path)
//   node ref_dump.js /path/to/dusp out_dir
// writes out_dir/<case>.f32 ([channel][n] little-endian) + out_dir/index.json; tests/golden/import_ref.py turns
// that into tests/golden/golden_ref.npz.
//
// Seeded noise: Math.random is replaced by the injected counter-based generator (splitmix64, identical to
// dusp_rand53 in include/dusp_b200.h): every Noise unit gets its own stream (seed, draw index).
const fs = require('fs'), path = require('path')
const [refDir, outDir] = process.argv.slice(2)
const dusp = require(path.join(refDir, 'src'))
const Noise = require(path.join(refDir, 'src/components/Noise'))

const M = (1n << 64n) - 1n
function rand53(seed, k) {
  let z = (BigInt(seed) + (BigInt(k) + 1n) * 0x9E3779B97F4A7C15n) & M
  z = ((z ^ (z >> 30n)) * 0xBF58476D1CE4E5B9n) & M
  z = ((z ^ (z >> 27n)) * 0x94D049BB133111EBn) & M
  z ^= z >> 31n
  return Number(z >> 11n) / 9007199254740992
}
// per-unit streams: draw 0 in the constructor (Noise.js:10), draws 1.. inside _tick (Noise.js:23)
let current = null
const realRandom = Math.random
Math.random = () => current ? rand53(current.seed, current.k++) : realRandom()
// stream seeds exactly as dusp_b200/config.py hands them out (host seed 12345 per case, Noise ordinal 0, 1, ..)
function mix64(z) {
  z &= M
  z = ((z ^ (z >> 30n)) * 0xBF58476D1CE4E5B9n) & M
  z = ((z ^ (z >> 27n)) * 0x94D049BB133111EBn) & M
  return z ^ (z >> 31n)
}
let hostSeed = 12345n, noiseOrdinal = 0n
function nextNoiseSeed() { return mix64((hostSeed ^ 0xD1B54A32D192ED03n) + (noiseOrdinal++) * 0x9E3779B97F4A7C15n) }
function seededNoise(f) {
  const st = { seed: nextNoiseSeed(), k: 0 }
  current = st
  const n = new Noise(f)
  current = null
  const tick = n._tick
  n._tick = function (clock) { current = st; try { return tick.call(this, clock) } finally { current = null } }
  return n
}

const C = dusp.components, P = dusp.patches
const cases = {  // the same circuits as tests/circuits.py (names match CASES there)
  osc440: [() => dusp.unDusp('[Osc 440]'), 1.0],
  readme_fm_literal: [() => dusp.unDusp('[Osc f:[Ramp from:200 to:100 duration:2]]'), 1.0],
  sum_pan: [() => dusp.unDusp('(O200 + O250) @ 0'), 1.0],
  w2_live: [() => { const o = new C.Osc(new C.Ramp(88200, 313.7, 313.7 / 2).trigger()); return new C.Multiply(o, o) }, 2.5],
  filter_lp: [() => new C.Filter(seededNoise(), 800, 'LP'), 0.5],
  w3_voice: [() => {
    const filt = new C.Filter(seededNoise(), 1234.5, 'LP')
    const mt = new P.MultiTapDelay(1, 1.0, filt)
    const tap = mt.addFeedback(0.137, 0.55, 0)
    return new C.Pan(new C.Sum(filt, tap), -0.3)
  }, 1.0],
}

;(async () => {
  fs.mkdirSync(outDir, { recursive: true })
  const index = {}
  for (const name in cases) {
    const [make, seconds] = cases[name]
    noiseOrdinal = 0n
    const cd = await dusp.renderChannelData(make(), seconds)
    const flat = new Float32Array(cd.length * cd[0].length)
    cd.forEach((c, i) => flat.set(c, i * c.length))
    fs.writeFileSync(path.join(outDir, name + '.f32'), Buffer.from(flat.buffer))
    index[name] = { channels: cd.length, samples: cd[0].length, seconds }
  }
  fs.writeFileSync(path.join(outDir, 'index.json'), JSON.stringify(index, null, 1))
})()
