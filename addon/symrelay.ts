// addon/symrelay.ts -- TypeScript glue (SOURCE ONLY; no Node.js in this image).
//
// Hooks the stream seam of the reference: every frame a connection writes or
// receives goes through the batcher instead of one synchronous libsodium call.
//   * `attach(peer)` is called from the hyperswarm 'connection' handler
//     (src/provider.ts:55-58): it adopts the connection's secret-stream states
//     (peer._encrypt / peer._decrypt, typed at src/types.ts:168-177) into a slot
//     and overrides `_write` so that NOTHING bypasses the GPU-owned nonce chain
//     (keep-alive frames included, SURVEY.md 7.3 item 4).
//   * the relay loop of src/provider.ts:240-257 stays as written: `peer.write(chunk)`
//     still returns a boolean and 'drain' still fires.
import { Peer } from "../src/types";
// eslint-disable-next-line @typescript-eslint/no-var-requires
const native = require("./symrelay.node");

const F_OPEN = 0x04000000;
const KIND: Record<string, number> = { ollama: 1, openwebui: 1, llamacpp: 2 }; // else 0 (utils.ts:45-50)

interface Pending { slot: number; data: Buffer; cb: (err?: Error | null) => void; peer: Peer }

export class SymRelayBatcher {
  private ctx: unknown;
  private arenaIn: Buffer;
  private arenaOut: Buffer;
  private queue: Pending[] = [];
  private timer: NodeJS.Timeout | null = null;

  constructor(gpuIds: number[], private flushMicros = 200, arenaBytes = 8 << 20) {
    this.ctx = native.create(Int32Array.from(gpuIds));
    this.arenaIn = Buffer.from(native.allocPinned(this.ctx, arenaBytes));
    this.arenaOut = Buffer.from(native.allocPinned(this.ctx, arenaBytes));
  }

  attach(peer: Peer, apiProvider: string): number {
    const enc = peer._encrypt, dec = peer._decrypt; // {key[32], state[52], header[24]} / {key, state, final}
    const slot: number = native.streamOpen(this.ctx, enc.key, enc.state, dec.key, dec.state, KIND[apiProvider] ?? 0);
    const self = this;
    // secret-stream `_write(data, cb)`: was alloc(len+20) + Push.next + rawStream.write
    (peer as any)._write = function (data: Buffer, cb: (err?: Error | null) => void) {
      self.queue.push({ slot, data, cb, peer });
      self.arm();
    };
    peer.once("close", () => native.streamClose(this.ctx, slot));
    return slot;
  }

  private arm() {
    if (this.queue.length >= 4096) return this.flush();
    if (!this.timer) this.timer = setTimeout(() => this.flush(), this.flushMicros / 1000);
  }

  private flush() {
    if (this.timer) { clearTimeout(this.timer); this.timer = null; }
    // at most one frame per slot per batch (MAC chain): later frames of a slot wait for the next batch
    const seen = new Set<number>(), batch: Pending[] = [], rest: Pending[] = [];
    for (const p of this.queue) (seen.has(p.slot) ? rest : (seen.add(p.slot), batch)).push(p);
    this.queue = rest;
    const desc = new Uint32Array(batch.length * 4), status = new Uint32Array(batch.length);
    let inOff = 0, outOff = 0;
    batch.forEach((p, i) => {
      p.data.copy(this.arenaIn, inOff);                       // the one host copy (V8 memory is not pinned)
      desc.set([p.slot, inOff, outOff + 12, p.data.length], i * 4);
      inOff += Math.max(16, (p.data.length + 15) & ~15);
      outOff += (12 + p.data.length + 20 + 15) & ~15;
    });
    const ticket = native.batchSubmit(this.ctx, desc, this.arenaIn.subarray(0, inOff), this.arenaOut.subarray(0, outOff), status);
    setImmediate(() => {                                      // production: threadsafe-function wake-up
      native.batchWait(this.ctx, ticket);
      batch.forEach((p, i) => {
        const o = desc[i * 4 + 2], wire = Buffer.from(this.arenaOut.subarray(o, o + p.data.length + 20));
        (p.peer as any).rawStream.write(wire);                // udx-native, unchanged
        p.cb(null);                                           // streamx: next write / 'drain'
      });
      if (this.queue.length) this.arm();
    });
  }
}
export { F_OPEN };
