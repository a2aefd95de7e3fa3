// addon/symrelay.ts -- TypeScript glue (SOURCE ONLY; no Node.js in this image).
//
// Hooks the stream seam of the reference: every frame a connection writes or receives goes through the
// library's per-frame front end instead of one synchronous libsodium call.
//   * `attach(peer, apiProvider)` is called from the hyperswarm 'connection' handler
//     (src/provider.ts:55-58).  It adopts the connection's secret-stream states (peer._encrypt /
//     peer._decrypt, typed at src/types.ts:168-177) into a slot and overrides BOTH directions:
//     `_write` (outbound: was alloc(len+20) + Push.next + rawStream.write) and `_onrawdata` (inbound: was
//     the uint24 reassembly + Pull.next + push), so that NOTHING bypasses the GPU-owned nonce chains --
//     keep-alive frames included (SURVEY.md 7.3 item 4).
//   * `write(peer, chunk, onDelta)` replaces the body of the relay loop, src/provider.ts:243-248: the SSE
//     `data:` test, JSON.parse and getChatDataFromProvider (src/utils.ts:16-52) run fused with the seal on
//     the GPU (SYMRELAY_F_EXTRACT); `onDelta(text)` is `completion += ...` of :243-246.  It returns what
//     `peer.write(chunk)` returns; 'drain' fires as before (:250-252).
//   * Nothing here blocks the event loop: the library's eventfd is watched by a uv_poll handle in the shim
//     (native.onReady) and completions are drained in its callback.  All frame memory is owned by the
//     library (chunks are copied into its pinned ring at submit, results are copied out at poll), so there
//     is no arena that a second flush could overwrite.
import { Peer } from "../src/types";
import { getChatDataFromProvider, safeParseStreamResponse } from "../src/utils";
// eslint-disable-next-line @typescript-eslint/no-var-requires
const native = require("./symrelay.node");

const F_EXTRACT = 0x08000000;
const EAGAIN = -11;
const ST = { OK: 0, BAD_MAC: 1, TOO_SHORT: 2, EXTRACT_THROW: 3, EXTRACT_UNSUPPORTED: 4, RETRY: 5, FAILED: 6 };
const KIND: Record<string, number> = { ollama: 1, openwebui: 1, llamacpp: 2 }; // else 0 (utils.ts:45-50)

interface PendingSeal { peer: Peer; chunk: Buffer; cb?: (err?: Error | null) => void; onDelta?: (s: string) => void; provider: string }
interface Conn { slot: number; provider: string; backlog: PendingSeal[]; rxBacklog: Buffer[]; wantDrain: boolean; held: PendingSeal | null }

export class SymRelayBatcher {
  private ctx: unknown;
  private seals = new Map<bigint | number, PendingSeal>();   // ticket -> what to do when it is sealed
  private opens = new Map<bigint | number, Peer>();          // ticket -> who gets the plaintext
  private conns = new Map<Peer, Conn>();
  private tickets = new BigInt64Array(256);

  constructor(gpuIds: number[], flushMicros = 100, maxStreams = 65536) {
    this.ctx = native.create(Int32Array.from(gpuIds), flushMicros, maxStreams);
    native.onReady(this.ctx, () => this.drain());             // eventfd -> uv_poll: called on the event loop
  }

  /** hyperswarm 'connection' (provider.ts:55-58), after secret-stream has exchanged its first messages. */
  attach(peer: Peer, apiProvider: string): number {
    const enc = (peer as any)._encrypt, dec = (peer as any)._decrypt; // {key[32], state[52], header[24]} / {key, state, final}
    const slot: number = native.streamOpen(this.ctx, enc.key, enc.state, dec.key, dec.state, KIND[apiProvider] ?? 0);
    if ((peer as any).keepAlive) native.setKeepAlive(this.ctx, slot, 1);
    const conn: Conn = { slot, provider: apiProvider, backlog: [], rxBacklog: [], wantDrain: false, held: null };
    this.conns.set(peer, conn);
    // outbound: every streamx write of this connection (relay chunks, JSON messages, keep-alives)
    (peer as any)._write = (data: Buffer, cb: (err?: Error | null) => void) => this.enqueue(conn, { peer, chunk: data, cb, provider: apiProvider });
    // inbound: raw transport bytes, cut anywhere (secret-stream `_onrawdata`, types.ts:156-159)
    (peer as any)._onrawdata = (data: Buffer) => this.feed(peer, conn, data);
    peer.once("close", () => this.detach(peer, conn));
    return slot;
  }

  /** The relay loop's body (provider.ts:243-248).  Returns false when the caller should await 'drain'. */
  write(peer: Peer, chunk: Buffer, onDelta: (text: string) => void): boolean {
    const conn = this.conns.get(peer)!;
    this.enqueue(conn, { peer, chunk, onDelta, provider: conn.provider });
    if (conn.backlog.length > 0) conn.wantDrain = true;       // ring full: `write() === false`
    return conn.backlog.length === 0;
  }

  /** One frame to the library; false when its ring is full (SYMRELAY_EAGAIN) or the connection is held.
   *  A frame that asks for extraction holds its connection until it has completed: the GPU may hand it back unsealed
   *  (EXTRACT_UNSUPPORTED) and it must then still reach the wire BEFORE the chunks that follow it.  One frame per
   *  stream per batch is the MAC chain's rule anyway; throughput comes from the number of connections. */
  private seal(conn: Conn, p: PendingSeal): boolean {
    if (conn.held !== null && conn.held !== p) return false;
    const t = native.submitSeal(this.ctx, conn.slot, p.chunk, p.onDelta ? F_EXTRACT : 0);
    if (t === EAGAIN) return false;
    if (p.onDelta) conn.held = p;
    this.seals.set(t, p);
    return true;
  }
  /** Frames of a connection go out in order: behind its backlog if there is one. */
  private enqueue(conn: Conn, p: PendingSeal) {
    if (conn.backlog.length > 0 || !this.seal(conn, p)) conn.backlog.push(p);
  }

  private feed(peer: Peer, conn: Conn, data: Buffer) {
    if (conn.rxBacklog.length > 0) { conn.rxBacklog.push(data); return; }
    let off = 0;
    while (off < data.length) {
      const r = native.rxFeed(this.ctx, conn.slot, data.subarray(off), this.tickets);   // throws on EPROTO -> destroy below
      for (let i = 0; i < Math.max(r.frames, 0); i++) this.opens.set(this.tickets[i], peer);
      off += r.consumed;
      if (r.frames === EAGAIN || (r.frames >= 0 && r.consumed === 0)) { conn.rxBacklog.push(data.subarray(off)); return; }
    }
  }

  /** Called by the shim when the library's eventfd fires: deliver everything that has finished. */
  private drain() {
    for (;;) {
      const done: any[] = native.poll(this.ctx, 1024);
      if (done.length === 0) break;
      for (const c of done) c.open ? this.opened(c) : this.sealed(c);
    }
    for (const [peer, conn] of this.conns) {                  // room again: backlogged frames in order, then 'drain'
      while (conn.backlog.length > 0 && this.seal(conn, conn.backlog[0])) conn.backlog.shift();
      const rx = conn.rxBacklog; conn.rxBacklog = [];
      for (const seg of rx) this.feed(peer, conn, seg);
      if (conn.wantDrain && conn.backlog.length === 0) { conn.wantDrain = false; peer.emit("drain"); }
    }
  }

  private sealed(c: any) {
    const p = this.seals.get(c.ticket)!; this.seals.delete(c.ticket);
    const conn = this.conns.get(p.peer);
    if (conn && conn.held === p && c.status !== ST.EXTRACT_UNSUPPORTED) conn.held = null;
    if (c.status === ST.OK) {
      (p.peer as any).rawStream.write(c.data);                // udx-native, unchanged: the wire frame incl. its uint24 prefix
      if (p.onDelta) p.onDelta(c.delta.toString());           // `completion += ...` (provider.ts:243-246)
      if (p.cb) p.cb(null);                                   // streamx: next write
    } else if (c.status === ST.EXTRACT_UNSUPPORTED && conn) {
      // a value shape the GPU extractor leaves to the reference's own helper (exponent numbers, arrays, deep nesting):
      // run it here, then seal the chunk without the extract flag
      try {
        const text = getChatDataFromProvider(p.provider, safeParseStreamResponse(p.chunk.toString()));
        const onDelta = p.onDelta; p.onDelta = undefined;     // second time round: seal only, still ahead of everything held back
        if (!this.seal(conn, p)) conn.backlog.unshift(p);
        conn.held = null;
        if (onDelta) onDelta(text as string);
      } catch (e) { p.peer.destroy(e as Error); }
    } else if (c.status === ST.EXTRACT_THROW) {
      // the reference throws a TypeError here (utils.ts:38-50) before peer.write: the stream is aborted, provider.ts:270-274
      const e = new TypeError("Cannot read properties of undefined");
      if (p.cb) p.cb(e); else p.peer.destroy(e);
    } else {
      const e = new Error(`symrelay: seal failed with status ${c.status}`);
      if (p.cb) p.cb(e); else p.peer.destroy(e);
    }
  }

  private opened(c: any) {
    const peer = this.opens.get(c.ticket)!; this.opens.delete(c.ticket);
    if (c.status === ST.OK) (peer as any).push(c.data);       // secret-stream `_incoming`: this.push(plain) -> peer.on('data'), provider.ts:174
    else if (c.status === ST.RETRY) { /* behind a REKEY-tagged frame: the library keeps such frames in order; resubmitted by rxFeed */ }
    else peer.destroy(new Error(c.status === ST.TOO_SHORT ? "Invalid message received" : "Could not decrypt message"));
  }

  private detach(peer: Peer, conn: Conn) {
    this.conns.delete(peer);
    const tryClose = () => { if (native.streamClose(this.ctx, conn.slot) === EAGAIN) setTimeout(tryClose, 1); };
    tryClose();                                               // frames of the slot may still be in a batch in flight
  }

  stats() { return native.schedStats(this.ctx); }
}
