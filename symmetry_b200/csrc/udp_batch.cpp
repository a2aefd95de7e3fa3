// udp_batch.cpp -- the step after the path: sealed frames to datagrams, many per system call.
//
// In the reference every `peer.write(chunk)` (/root/reference/src/provider.ts:248) ends, below secret-stream, in udx-native
// 1.10.3 (/root/reference/package-lock.json:6241-6253; the stream/socket shapes are typed at src/types.ts:91-122) handing
// ONE packet to the kernel per frame.  Once the crypto is batched that system call is the host's next per-frame cost.
// This is the send primitive a batched transport would call with the completions symrelay_poll delivers: every wire frame
// is cut into datagrams of at most `mtu` bytes (16-byte header + payload slice, gathered straight from the library's pinned
// ring -- no copy) and up to 1,024 datagrams go out per sendmmsg(2).  It does not speak UDX (acknowledgements,
// retransmission and congestion control stay in udx-native); the header only says where a slice belongs in its
// connection's byte stream so that a receiver can put the stream back together for symrelay_rx_feed.
// Host-only code: no CUDA in this file.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <errno.h>
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/types.h>

#include <vector>

#include "../../include/symrelay.h"

extern "C" int symrelay_udp_send_batch(int fd, const symrelay_completion *comp, int n, const void *const *dst, const uint32_t *dst_len,
                                       const uint32_t *conn_id, uint64_t *stream_off, uint32_t mtu, uint32_t flags,
                                       symrelay_udp_stats *stats) {
    if (fd < 0 || n < 0 || (n && (!comp || !conn_id || !stream_off)) || mtu < SYMRELAY_UDP_HDR + 1u || mtu > 65507u) return SYMRELAY_EINVAL;
    const uint32_t room = mtu - SYMRELAY_UDP_HDR;
    const size_t per_call = (flags & SYMRELAY_UDP_ONE_PER_CALL) ? 1u : 1024u;   // the flag reproduces one sendmsg per datagram
    static thread_local std::vector<mmsghdr> msgs;
    static thread_local std::vector<iovec> iov;
    static thread_local std::vector<symrelay_udp_hdr> hdrs;
    size_t total = 0;
    for (int i = 0; i < n; i++)
        if (SYMRELAY_STATUS_CODE(comp[i].status) == SYMRELAY_ST_OK && !(comp[i].flags & SYMRELAY_F_OPEN) && comp[i].len)
            total += (comp[i].len + room - 1u) / room;
    msgs.resize(total); iov.resize(2 * total); hdrs.resize(total);
    size_t k = 0;
    uint64_t frames = 0, bytes = 0;
    for (int i = 0; i < n; i++) {
        const symrelay_completion &c = comp[i];
        if (SYMRELAY_STATUS_CODE(c.status) != SYMRELAY_ST_OK || (c.flags & SYMRELAY_F_OPEN) || !c.len) continue;   // only sealed frames travel
        frames++;
        for (uint32_t off = 0; off < c.len; off += room, k++) {
            const uint32_t len = c.len - off < room ? c.len - off : room;
            hdrs[k].magic = SYMRELAY_UDP_MAGIC;
            hdrs[k].conn = conn_id[i];
            hdrs[k].offset = stream_off[i] + off;
            iov[2 * k].iov_base = &hdrs[k]; iov[2 * k].iov_len = SYMRELAY_UDP_HDR;
            iov[2 * k + 1].iov_base = const_cast<uint8_t *>(c.data) + off; iov[2 * k + 1].iov_len = len;
            memset(&msgs[k], 0, sizeof(mmsghdr));
            msgs[k].msg_hdr.msg_iov = &iov[2 * k];
            msgs[k].msg_hdr.msg_iovlen = 2;
            if (dst && dst[i]) { msgs[k].msg_hdr.msg_name = const_cast<void *>(dst[i]); msgs[k].msg_hdr.msg_namelen = dst_len[i]; }
            bytes += len;
        }
        stream_off[i] += c.len;
    }
    uint64_t calls = 0;
    for (size_t at = 0; at < total;) {
        const size_t want = total - at < per_call ? total - at : per_call;
        const int sent = sendmmsg(fd, &msgs[at], (unsigned)want, 0);
        if (sent < 0) {
            if (errno == EINTR) continue;
            if (stats) { stats->frames += frames; stats->datagrams += at; stats->syscalls += calls; stats->payload_bytes += bytes; }
            return errno == EAGAIN || errno == EWOULDBLOCK || errno == ENOBUFS ? SYMRELAY_EAGAIN : SYMRELAY_EINVAL;
        }
        calls++;
        at += (size_t)sent;
    }
    if (stats) { stats->frames += frames; stats->datagrams += total; stats->syscalls += calls; stats->payload_bytes += bytes; }
    return (int)total;
}
