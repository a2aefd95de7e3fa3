#!/usr/bin/env node
// tests/golden/check_extract_cases.js -- pins the delta-extraction oracle against the REFERENCE's own code.
//
// Nothing in this repository's image can run JavaScript (no node, deno, bun, quickjs: SURVEY.md 8(c)), so
// tests/golden/extract_cases.json was produced by oracle/extract.py, a Python restatement of
// src/utils.ts:16-52 + src/provider.ts:243-246, and parity for rows A2-A4 is "unpinned" until this script has
// been run once on a machine with Node.  It replays every case through the reference's compiled utils:
//
//     cd <symmetry checkout> && npm ci && npm run build        # tsc -> dist/utils.js, dist/constants.js
//     node <this repo>/tests/golden/check_extract_cases.js <symmetry checkout>/dist
//
// Per case {provider, chunk (hex), status, delta (hex)} it does what the relay loop does with one chunk:
//     completion += getChatDataFromProvider(provider, safeParseStreamResponse(chunk.toString()))
// status 0: no exception, delta = the string `+=` appended, as WTF-8 (lone surrogates kept: JSON.parse
//           accepts "\ud800"; Buffer.from(s, 'utf8') would turn it into U+FFFD, so the encoding is done by hand);
// status 1: the call threw a TypeError (the stream is aborted, provider.ts:270-274).
// Exit code 0 and "994/994 cases agree" pin the oracle; any difference is printed with the case index.
'use strict'
const fs = require('fs')
const path = require('path')

const dist = process.argv[2]
if (!dist) { console.error('usage: check_extract_cases.js <path to the reference\'s dist/ directory>'); process.exit(2) }
const { getChatDataFromProvider, safeParseStreamResponse } = require(path.resolve(dist, 'utils.js'))

function wtf8 (s) {            // UTF-16 code units -> bytes; a surrogate without its partner becomes a 3-byte sequence
  const out = []
  for (let i = 0; i < s.length; i++) {
    let c = s.charCodeAt(i)
    if (c >= 0xd800 && c <= 0xdbff && i + 1 < s.length) {
      const d = s.charCodeAt(i + 1)
      if (d >= 0xdc00 && d <= 0xdfff) { c = 0x10000 + ((c - 0xd800) << 10) + (d - 0xdc00); i++ }
    }
    if (c < 0x80) out.push(c)
    else if (c < 0x800) out.push(0xc0 | (c >> 6), 0x80 | (c & 63))
    else if (c < 0x10000) out.push(0xe0 | (c >> 12), 0x80 | ((c >> 6) & 63), 0x80 | (c & 63))
    else out.push(0xf0 | (c >> 18), 0x80 | ((c >> 12) & 63), 0x80 | ((c >> 6) & 63), 0x80 | (c & 63))
  }
  return Buffer.from(out).toString('hex')
}

const file = path.join(__dirname, 'extract_cases.json')
const cases = JSON.parse(fs.readFileSync(file, 'utf8')).cases
let bad = 0
cases.forEach((c, i) => {
  const chunk = Buffer.from(c.chunk, 'hex')
  let status = 0
  let delta = ''
  try {
    let completion = ''
    completion += getChatDataFromProvider(c.provider, safeParseStreamResponse(chunk.toString()))   // provider.ts:243-246
    delta = wtf8(completion)
  } catch (e) {
    if (!(e instanceof TypeError)) throw e
    status = 1
  }
  if (status !== c.status || (status === 0 && delta !== c.delta)) {
    bad++
    console.log(`case ${i} (${c.provider}) differs: reference status ${status} delta ${delta} | fixture status ${c.status} delta ${c.delta} | chunk ${JSON.stringify(chunk.toString())}`)
  }
})
console.log(`${cases.length - bad}/${cases.length} cases agree with the reference's utils.js`)
process.exit(bad ? 1 : 0)
