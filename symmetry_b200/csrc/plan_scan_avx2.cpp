// plan_scan_avx2.cpp -- pass 1 of the planner (plan.h: symrelay_plan_scan_scalar) eight descriptors at a
// time.  The library is built for baseline x86-64; this one function carries its own target attribute
// and plan.h calls it only when the CPU reports AVX2.  Same results as the scalar loop, element for
// element (tests/test_abi_symbols.py compares the two on random batches).
//
// Why: through the host path the planner runs once per batch on the caller's thread; the compiler's
// own vectorisation of the scalar loop (16-byte vectors, array-of-structs loads through pinsr/pextr)
// took 4.1 us of an 11.4 us batch of 2,048 descriptors.
#include <stdint.h>

#include "../../include/symrelay.h"

#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>

#define AVX2 __attribute__((target("avx2")))

// a > b, unsigned
AVX2 static inline __m256i gt_u32(__m256i a, __m256i b) {
    return _mm256_xor_si256(_mm256_cmpeq_epi32(_mm256_max_epu32(b, a), b), _mm256_set1_epi32(-1));
}

extern "C" AVX2 __attribute__((visibility("hidden"))) void symrelay_plan_scan_avx2(
    const symrelay_desc *desc, uint32_t n, uint32_t *nc, uint32_t *row_lo, uint32_t *row_hi, uint32_t in_bytes,
    uint32_t out_bytes, uint32_t max_streams, uint32_t out[6] /* bad, oob, any_flags, nc_min, nc_max, overlap */) {
    const __m256i v_len_mask = _mm256_set1_epi32((int)SYMRELAY_LEN_MASK), v_one = _mm256_set1_epi32(1);
    const __m256i v_max_frame = _mm256_set1_epi32((int)SYMRELAY_MAX_FRAME), v_15 = _mm256_set1_epi32(15);
    const __m256i v_12 = _mm256_set1_epi32(12), v_20 = _mm256_set1_epi32(20), v_63 = _mm256_set1_epi32(63);
    const __m256i v_2 = _mm256_set1_epi32(2), v_in = _mm256_set1_epi32((int)in_bytes), v_out = _mm256_set1_epi32((int)out_bytes);
    const __m256i v_streams_m1 = _mm256_set1_epi32((int)(max_streams - 1u));
    const __m256i order = _mm256_setr_epi32(0, 4, 1, 5, 2, 6, 3, 7);   // [0,2,4,6,1,3,5,7] -> [0..7]
    __m256i v_bad = _mm256_setzero_si256(), v_oob = v_bad, v_flags = v_bad, v_min = _mm256_set1_epi32(-1), v_max = v_bad;
    uint32_t i = 0;
    for (; i + 8u <= n; i += 8u) {
        const __m256i *p = reinterpret_cast<const __m256i *>(desc + i);
        const __m256i r0 = _mm256_loadu_si256(p), r1 = _mm256_loadu_si256(p + 1), r2 = _mm256_loadu_si256(p + 2),
                      r3 = _mm256_loadu_si256(p + 3);
        // {slot, in_off, out_off, len_flags} x 8 -> four vectors, elements in the order 0,2,4,6,1,3,5,7
        const __m256i t0 = _mm256_unpacklo_epi32(r0, r1), t1 = _mm256_unpackhi_epi32(r0, r1);
        const __m256i t2 = _mm256_unpacklo_epi32(r2, r3), t3 = _mm256_unpackhi_epi32(r2, r3);
        const __m256i slot = _mm256_unpacklo_epi64(t0, t2), in_off = _mm256_unpackhi_epi64(t0, t2);
        const __m256i out_off = _mm256_unpacklo_epi64(t1, t3), lf = _mm256_unpackhi_epi64(t1, t3);

        const __m256i mlen = _mm256_and_si256(lf, v_len_mask);
        const __m256i op = _mm256_and_si256(_mm256_srli_epi32(lf, 26), v_one);          // SYMRELAY_F_OPEN
        const __m256i c = _mm256_add_epi32(v_2, _mm256_srli_epi32(_mm256_add_epi32(mlen, v_63), 6));
        const __m256i op20 = _mm256_mullo_epi32(op, v_20), op12 = _mm256_mullo_epi32(op, v_12);
        const __m256i need_in = _mm256_add_epi32(mlen, op20);
        const __m256i need_out = _mm256_sub_epi32(_mm256_add_epi32(mlen, v_20), op20);
        const __m256i ones = _mm256_set1_epi32(-1);
        __m256i bad = gt_u32(mlen, v_max_frame);
        bad = _mm256_or_si256(bad, _mm256_xor_si256(_mm256_cmpeq_epi32(_mm256_and_si256(in_off, v_15), op12), ones));
        bad = _mm256_or_si256(bad, _mm256_xor_si256(
                                       _mm256_cmpeq_epi32(_mm256_and_si256(out_off, v_15), _mm256_sub_epi32(v_12, op12)), ones));
        v_bad = _mm256_or_si256(v_bad, bad);
        __m256i oob = gt_u32(need_in, v_in);
        oob = _mm256_or_si256(oob, gt_u32(in_off, _mm256_sub_epi32(v_in, need_in)));
        oob = _mm256_or_si256(oob, gt_u32(need_out, v_out));
        oob = _mm256_or_si256(oob, gt_u32(out_off, _mm256_sub_epi32(v_out, need_out)));
        oob = _mm256_or_si256(oob, gt_u32(slot, v_streams_m1));                          // slot >= max_streams
        v_oob = _mm256_or_si256(v_oob, oob);
        v_flags = _mm256_or_si256(v_flags, lf);
        v_min = _mm256_min_epu32(v_min, c);
        v_max = _mm256_max_epu32(v_max, c);
        const __m256i lo = _mm256_srli_epi32(out_off, 4);
        const __m256i hi = _mm256_add_epi32(
            lo, _mm256_srli_epi32(_mm256_add_epi32(_mm256_add_epi32(_mm256_and_si256(out_off, v_15), need_out), v_15), 4));
        _mm256_storeu_si256(reinterpret_cast<__m256i *>(nc + i), _mm256_permutevar8x32_epi32(c, order));
        _mm256_storeu_si256(reinterpret_cast<__m256i *>(row_lo + i), _mm256_permutevar8x32_epi32(lo, order));
        _mm256_storeu_si256(reinterpret_cast<__m256i *>(row_hi + i), _mm256_permutevar8x32_epi32(hi, order));
    }
    uint32_t tmp[8];
    uint32_t bad = 0, oob = 0, any_flags = 0, nc_min = 0xffffffffu, nc_max = 0;
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_bad);
    for (int k = 0; k < 8; k++) bad |= tmp[k];
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_oob);
    for (int k = 0; k < 8; k++) oob |= tmp[k];
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_flags);
    for (int k = 0; k < 8; k++) any_flags |= tmp[k];
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_min);
    for (int k = 0; k < 8; k++) nc_min = tmp[k] < nc_min ? tmp[k] : nc_min;
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_max);
    for (int k = 0; k < 8; k++) nc_max = tmp[k] > nc_max ? tmp[k] : nc_max;
    bad = bad ? 1u : 0u;
    oob = oob ? 1u : 0u;
    if (max_streams == 0u && n >= 8u) oob = 1u;   // every slot is out of range; the vector test compared against 2^32-1
    for (; i < n; i++) {   // the last n % 8 descriptors: the scalar loop of plan.h
        const uint32_t lf = desc[i].len_flags, in_off = desc[i].in_off, out_off = desc[i].out_off;
        const uint32_t mlen = lf & SYMRELAY_LEN_MASK;
        const uint32_t op = (lf >> 26) & 1u;
        const uint32_t c = 2u + ((mlen + 63u) >> 6);
        const uint32_t need_in = mlen + 20u * op, need_out = mlen + 20u - 20u * op;
        bad |= (uint32_t)(mlen > SYMRELAY_MAX_FRAME) | (uint32_t)((in_off & 15u) != 12u * op) |
               (uint32_t)((out_off & 15u) != 12u - 12u * op);
        oob |= (uint32_t)(need_in > in_bytes) | (uint32_t)(in_off > in_bytes - need_in) | (uint32_t)(need_out > out_bytes) |
               (uint32_t)(out_off > out_bytes - need_out) | (uint32_t)(desc[i].slot >= max_streams);
        any_flags |= lf;
        nc[i] = c;
        nc_min = c < nc_min ? c : nc_min;
        nc_max = c > nc_max ? c : nc_max;
        row_lo[i] = out_off >> 4;
        row_hi[i] = (out_off >> 4) + (((out_off & 15u) + need_out + 15u) >> 4);
    }
    // output rows strictly increasing in descriptor order?
    uint32_t overlap = 0;
    uint32_t j = 1;
    __m256i v_ov = _mm256_setzero_si256();
    for (; j + 8u <= n; j += 8u) {
        const __m256i lo = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(row_lo + j));
        const __m256i hi = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(row_hi + j - 1));
        v_ov = _mm256_or_si256(v_ov, gt_u32(hi, lo));   // row_lo[j] < row_hi[j-1]
    }
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(tmp), v_ov);
    for (int k = 0; k < 8; k++) overlap |= tmp[k];
    overlap = overlap ? 1u : 0u;
    for (; j < n; j++) overlap |= (uint32_t)(row_lo[j] < row_hi[j - 1]);
    out[0] = bad; out[1] = oob; out[2] = any_flags; out[3] = nc_min; out[4] = nc_max; out[5] = overlap;
}

extern "C" __attribute__((visibility("hidden"))) int symrelay_plan_scan_has_avx2(void) {
    __builtin_cpu_init();
    return __builtin_cpu_supports("avx2") ? 1 : 0;
}
#else
extern "C" void symrelay_plan_scan_avx2(const symrelay_desc *, uint32_t, uint32_t *, uint32_t *, uint32_t *, uint32_t, uint32_t,
                                        uint32_t, uint32_t *) {}
extern "C" int symrelay_plan_scan_has_avx2(void) { return 0; }
#endif
