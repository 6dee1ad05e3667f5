/*
 * pint_oracle.c -- CPU restatement of pint's per-element algorithms.
 * TEST INFRASTRUCTURE ONLY (see pint_oracle.h).  Parity status: PINNED
 * (tests/test_oracle_golden.py).  "pint.hpp:N" = /root/reference/include/pint/pint.hpp:N.
 */
#include "pint_oracle.h"

#include <string.h>

static inline uint64_t ones64(uint32_t w)
{
    /* all_ones<T, Bits>, pint.hpp:340-348 (the 64 case is the specialisation
     * that avoids the undefined 1 << 64). */
    return w >= 64 ? ~0ULL : ((1ULL << w) - 1ULL);
}

static inline int popcount64(uint64_t x)
{
    /* bit_count, pint.hpp:351-353 */
    int c = 0;
    while (x) { x &= x - 1; ++c; }
    return c;
}

uint32_t pint_oracle_storage_bytes(uint32_t total_bits)
{
    /* make_packed_int, pint.hpp:789-795: (clp2(sum - 1) + 7) & ~7 selects
     * 8/16/32/64; anything else hits the static_assert at :727-728. */
    if (total_bits == 0 || total_bits > 64) return 0;
    uint64_t x = (uint64_t)total_bits - 1; /* clp2<NumberMinus1>, pint.hpp:710-722 */
    x |= x >> 32; x |= x >> 16; x |= x >> 8; x |= x >> 4; x |= x >> 2; x |= x >> 1;
    uint64_t rounded = ((x + 1) + 7) & ~7ULL;
    switch (rounded) {
    case 8: return 1;
    case 16: return 2;
    case 32: return 4;
    case 64: return 8;
    default: return 0;
    }
}

int pint_oracle_validate(const pint_oracle_layout* l)
{
    if (!l) return -1;
    if (l->elem_bytes != 1 && l->elem_bytes != 2 && l->elem_bytes != 4 && l->elem_bytes != 8)
        return -2; /* pint.hpp:731-734, 761-762 */
    if (l->nfields < 1 || l->nfields > 64) return -3;
    uint32_t total = 0;
    for (uint32_t i = 0; i < l->nfields; ++i) {
        if (l->bits[i] == 0) return -4; /* Bits - 1 shifts (pint.hpp:327, 463) need Bits >= 1 */
        total += l->bits[i];
    }
    if (total > 8 * l->elem_bytes) return -5; /* pint.hpp:763-764 */
    return 0;
}

int pint_oracle_make_masks(const pint_oracle_layout* l, pint_oracle_masks* m)
{
    int rc = pint_oracle_validate(l);
    if (rc) return rc;
    memset(m, 0, sizeof(*m));
    m->word = ones64(8 * l->elem_bytes);
    uint32_t off = 0, maxb = 0;
    int all_same = 1;
    for (uint32_t i = 0; i < l->nfields; ++i) {
        uint32_t w = l->bits[i];
        m->off[i] = (uint8_t)off;              /* mask_offsets_vector, pint.hpp:288-292 */
        m->L |= 1ULL << off;                   /* mask_loorder, pint.hpp:332-337 */
        m->H |= 1ULL << (off + w - 1);         /* mask_hiorder, pint.hpp:324-329 */
        if (w > maxb) maxb = w;                /* find_max, pint.hpp:277-285 */
        if (w != l->bits[0]) all_same = 0;     /* all_same, pint.hpp:225-231 */
        /* unique<> keeps first-appearance order, pint.hpp:97-110; the per-width
         * LSB masks are unsigned_saturation_mask_type_2, pint.hpp:503-515 */
        uint32_t k = 0;
        while (k < m->nuniq && m->uniq_w[k] != w) ++k;
        if (k == m->nuniq) { m->uniq_w[k] = (uint8_t)w; m->nuniq++; }
        m->uniq_lo[k] |= 1ULL << off;
        off += w;
    }
    m->total = off;
    m->maxbits = maxb;
    m->A = ones64(off);
    /* detect_saturation_mask_type, pint.hpp:443-456 with
     * is_saturation_mask_of_type_1, pint.hpp:424-441 */
    if (all_same) {
        m->sat_type = 0;
    } else {
        uint32_t cnt = 0;
        for (uint32_t k = 0; k < m->nuniq; ++k)
            cnt += (uint32_t)popcount64((m->H >> (m->uniq_w[k] - 1)) & m->L);
        m->sat_type = (cnt == l->nfields) ? 1 : 2;
    }
    return 0;
}

/* dispatch_make_unsigned_saturation_mask: move every field's MSB flag to that
 * field's LSB.  type 0 pint.hpp:459-464, type 1 :467-490, type 2 :493-542. */
static inline uint64_t lsbflags(const pint_oracle_masks* m, uint64_t c)
{
    uint64_t r = 0;
    switch (m->sat_type) {
    case 0:
        return c >> (m->uniq_w[0] - 1);
    case 1:
        for (uint32_t k = 0; k < m->nuniq; ++k) r |= c >> (m->uniq_w[k] - 1);
        return r & m->L;
    default:
        for (uint32_t k = 0; k < m->nuniq; ++k) r |= (c >> (m->uniq_w[k] - 1)) & m->uniq_lo[k];
        return r;
    }
}

/* make_unsigned_saturation_mask, pint.hpp:544-551 */
static inline uint64_t usat_mask(const pint_oracle_masks* m, uint64_t c)
{
    return ((c << 1) - lsbflags(m, c)) & m->word;
}

/* make_signed_saturation_mask, pint.hpp:563-567 */
static inline uint64_t ssat_mask(const pint_oracle_masks* m, uint64_t ov)
{
    return ov - lsbflags(m, ov);
}

/* apply_signed_saturation, pint.hpp:569-574 */
static inline uint64_t apply_ssat(const pint_oracle_masks* m, uint64_t sum, uint64_t ov)
{
    return ((sum ^ ov) | ssat_mask(m, ov)) ^ ssat_mask(m, ov & ~sum);
}

/* carry_add_vector pint.hpp:375-378, carry_sub_vector :380-383,
 * overflow_signed_sub_vector :385-388 */
static inline uint64_t carry_add(uint64_t a, uint64_t b) { return (a & b) | ((a | b) & ~(a + b)); }
static inline uint64_t carry_sub(uint64_t a, uint64_t b) { return (~a & b) | ((~(a ^ b)) & (a - b)); }
static inline uint64_t ovf_ssub(uint64_t a, uint64_t b, uint64_t r) { return (~a & b & r) | (a & ~(b | r)); }

/* add_wrap, pint.hpp:826-838 */
static inline uint64_t add_wrap(const pint_oracle_masks* m, uint64_t a, uint64_t b)
{
    const uint64_t M1 = ~m->H & m->A;
    return ((a & M1) + (b & M1)) ^ ((a ^ b) & m->H);
}

/* sub_wrap, pint.hpp:870-884 */
static inline uint64_t sub_wrap(const pint_oracle_masks* m, uint64_t a, uint64_t b)
{
    const uint64_t M1 = ~m->H & m->A;
    return (((a & M1) + (~b & M1) + (m->L & M1)) ^ ((a ^ ~b) & m->H) ^ (m->H & m->L));
}

/* interleave, pint.hpp:603-606 */
static inline uint64_t interleave(uint64_t a, uint64_t b, uint64_t mask) { return (a & mask) | (b & ~mask); }

static inline uint64_t binary1(int op, const pint_oracle_masks* m, uint64_t a, uint64_t b)
{
    const uint64_t H = m->H;
    uint64_t r;
    switch (op) {
    case PINT_O_ADD_WRAP:
        r = add_wrap(m, a, b);
        break;
    case PINT_O_ADD_UNSIGNED_SATURATE: /* pint.hpp:840-855 + detail :554-558 */
        r = add_wrap(m, a, b) | usat_mask(m, carry_add(a, b) & H);
        break;
    case PINT_O_ADD_SIGNED_SATURATE: { /* pint.hpp:857-866 + detail :576-582 */
        uint64_t sum = add_wrap(m, a, b);
        r = apply_ssat(m, sum, (~(a ^ b)) & (sum ^ b) & H);
        break;
    }
    case PINT_O_SUB_WRAP:
        r = sub_wrap(m, a, b);
        break;
    case PINT_O_SUB_UNSIGNED_SATURATE: { /* pint.hpp:886-908 */
        uint64_t nb = ~b & m->word; /* packed(~b.value()) truncates to Integer, :768 */
        uint64_t t = add_wrap(m, a, nb) | usat_mask(m, carry_sub(a, b) & H);
        r = add_wrap(m, t, m->L);
        break;
    }
    case PINT_O_SUB_SIGNED_SATURATE: { /* pint.hpp:910-919 + detail :584-590 */
        uint64_t diff = sub_wrap(m, a, b) & m->word;
        r = apply_ssat(m, diff, ovf_ssub(a, b, diff) & H);
        break;
    }
    case PINT_O_OR: r = a | b; break;   /* pint.hpp:781 */
    case PINT_O_AND: r = a & b; break;  /* pint.hpp:782 */
    case PINT_O_XOR: r = a ^ b; break;  /* pint.hpp:783 */
    case PINT_O_MIN_UNSIGNED: /* pint.hpp:936-950 */
        r = interleave(a, b, usat_mask(m, carry_sub(a, b) & H));
        break;
    case PINT_O_MAX_UNSIGNED: /* pint.hpp:952-966 */
        r = interleave(a, b, usat_mask(m, carry_sub(b, a) & H));
        break;
    case PINT_O_MIN_SIGNED: /* pint.hpp:968-985 */
        r = interleave(a, b, usat_mask(m, carry_sub(a ^ H, b ^ H) & H));
        break;
    case PINT_O_MAX_SIGNED: /* pint.hpp:987-1004 */
        r = interleave(a, b, usat_mask(m, carry_sub(b ^ H, a ^ H) & H));
        break;
    default:
        r = 0;
    }
    return r & m->word;
}

uint64_t pint_oracle_binary1(int op, const pint_oracle_masks* m, uint64_t a, uint64_t b)
{
    return binary1(op, m, a & m->word, b & m->word);
}

/* shift_left, pint.hpp:1006-1017 with detail::shift_left :654-668 and
 * shift_left_mask :630-651.  For amount >= max(Bits) the reference's outer
 * guard ((sign_bit(max - amount - 1) - 1) & ...) yields 0. */
static inline uint64_t shift_left1(const pint_oracle_masks* m, uint64_t v, uint64_t amount)
{
    if (amount >= m->maxbits) return 0;
    uint64_t mask = 0;
    if (m->nuniq == 1) {
        /* pint.hpp:667.  A single full-width field (<uint32_t,32>, <uint64_t,64>)
         * with amount == 0 makes the reference shift by the type's width, which
         * is undefined behaviour there; the port takes the arithmetic reading
         * (2^W mod 2^W == 0, so the mask is all ones and the value is returned).
         * Tests exclude that one case from reference comparisons. */
        const uint64_t s = m->uniq_w[0] - amount;
        mask = (s >= 64 ? 0 : (m->L << s)) - m->L;
    } else {
        for (uint32_t k = 0; k < m->nuniq; ++k) {
            /* saturate_to_zero(bits - amount), pint.hpp:617-620 */
            uint64_t s = m->uniq_w[k] > amount ? m->uniq_w[k] - amount : 0;
            mask |= (m->uniq_lo[k] << s) - m->uniq_lo[k]; /* pint.hpp:634 */
        }
    }
    return ((mask & v) << amount) & m->word;
}

/* shift_right_unsigned, pint.hpp:1019-1029 with detail :692-705 and
 * shift_right_mask :670-690. */
static inline uint64_t shift_right1(const pint_oracle_masks* m, uint64_t v, uint64_t amount)
{
    if (amount >= m->maxbits) return 0;
    uint64_t mask = 0;
    if (m->nuniq == 1) {
        mask = (m->L << amount) - m->L; /* pint.hpp:704 */
    } else {
        for (uint32_t k = 0; k < m->nuniq; ++k) {
            /* find_min(bits, amount), pint.hpp:622-625 */
            uint64_t s = m->uniq_w[k] < amount ? m->uniq_w[k] : amount;
            mask |= (m->uniq_lo[k] << s) - m->uniq_lo[k]; /* pint.hpp:674 */
        }
    }
    return ((~mask & v) >> amount) & m->word;
}

uint64_t pint_oracle_shift_left1(const pint_oracle_masks* m, uint64_t v, uint64_t amount)
{
    return shift_left1(m, v & m->word, amount);
}

uint64_t pint_oracle_shift_right_unsigned1(const pint_oracle_masks* m, uint64_t v, uint64_t amount)
{
    return shift_right1(m, v & m->word, amount);
}

/* get<Index>, pint.hpp:799-807 */
static inline uint64_t get1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t i, uint64_t v)
{
    return (v >> m->off[i]) & ones64(l->bits[i]);
}

/* get_signed<Index>, pint.hpp:809-822: shift the field's MSB to the Integer's
 * MSB, reinterpret as make_signed<Integer>, arithmetic shift back. */
static inline int64_t get_signed1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t i, uint64_t v)
{
    const uint32_t W = 8 * l->elem_bytes;
    const uint32_t hi = m->off[i] + l->bits[i];
    uint64_t x = (v << (W - hi)) & m->word;
    /* sign-extend from W bits to 64 (the static_cast to the signed Integer) */
    int64_t sx = (int64_t)(x << (64 - W)) >> (64 - W);
    return sx >> (W - l->bits[i]);
}

/* slice<Start, End>, pint.hpp:923-932 */
static inline uint64_t slice1(const pint_oracle_layout* l, uint32_t s, uint32_t e, uint64_t v)
{
    uint32_t lo = 0, mid = 0;
    for (uint32_t i = 0; i < s; ++i) lo += l->bits[i];
    for (uint32_t i = s; i < e; ++i) mid += l->bits[i];
    return (v >> lo) & ones64(mid);
}

uint64_t pint_oracle_get1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t index, uint64_t v)
{
    return get1(l, m, index, v & m->word);
}

int64_t pint_oracle_get_signed1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t index, uint64_t v)
{
    return get_signed1(l, m, index, v & m->word);
}

uint64_t pint_oracle_slice1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t start, uint32_t end, uint64_t v)
{
    return slice1(l, start, end, v & m->word);
}

/* field constructor -> make_truncate -> make_truncated_int,
 * pint.hpp:770-774, 592-601, 390-407 */
uint64_t pint_oracle_make_truncate1(const pint_oracle_layout* l, const pint_oracle_masks* m, const uint64_t* values)
{
    uint64_t r = 0;
    for (uint32_t i = 0; i < l->nfields; ++i)
        r |= (values[i] & ones64(l->bits[i])) << m->off[i];
    return r & m->word;
}

/* ------------------------------------------------------------------ */
/* Independent naive per-field model (not derived from the SWAR code). */

typedef __int128 i128;

static i128 field_signed(uint64_t v, uint32_t w)
{
    i128 x = (i128)v;
    if (w < 128 && ((v >> (w - 1)) & 1)) x -= ((i128)1 << w);
    return x;
}

uint64_t pint_oracle_naive_binary1(int op, const pint_oracle_layout* l, uint64_t a, uint64_t b)
{
    const uint64_t word = ones64(8 * l->elem_bytes);
    a &= word; b &= word;
    if (op == PINT_O_OR) return a | b;
    if (op == PINT_O_AND) return a & b;
    if (op == PINT_O_XOR) return a ^ b;
    uint64_t r = 0;
    uint32_t off = 0;
    for (uint32_t i = 0; i < l->nfields; ++i) {
        const uint32_t w = l->bits[i];
        const uint64_t fm = ones64(w);
        const uint64_t ua = (a >> off) & fm, ub = (b >> off) & fm;
        const i128 sa = field_signed(ua, w), sb = field_signed(ub, w);
        const i128 smax = ((i128)1 << (w - 1)) - 1, smin = -((i128)1 << (w - 1));
        i128 t;
        uint64_t f;
        switch (op) {
        case PINT_O_ADD_WRAP: f = (uint64_t)((i128)ua + (i128)ub) & fm; break;
        case PINT_O_ADD_UNSIGNED_SATURATE:
            t = (i128)ua + (i128)ub; f = t > (i128)fm ? fm : (uint64_t)t; break;
        case PINT_O_ADD_SIGNED_SATURATE:
            t = sa + sb; t = t > smax ? smax : (t < smin ? smin : t); f = (uint64_t)t & fm; break;
        case PINT_O_SUB_WRAP: f = (uint64_t)((i128)ua - (i128)ub) & fm; break;
        case PINT_O_SUB_UNSIGNED_SATURATE: f = ua > ub ? ua - ub : 0; break;
        case PINT_O_SUB_SIGNED_SATURATE:
            t = sa - sb; t = t > smax ? smax : (t < smin ? smin : t); f = (uint64_t)t & fm; break;
        case PINT_O_MIN_UNSIGNED: f = ua < ub ? ua : ub; break;
        case PINT_O_MAX_UNSIGNED: f = ua > ub ? ua : ub; break;
        case PINT_O_MIN_SIGNED: f = sa < sb ? ua : ub; break;
        case PINT_O_MAX_SIGNED: f = sa > sb ? ua : ub; break;
        default: f = 0;
        }
        r |= f << off;
        off += w;
    }
    /* bits above sum(Bits): zero for add/sub (M1 and H exclude them), taken
     * from b for min/max (interleave with a mask confined to the fields). */
    if (op >= PINT_O_MIN_UNSIGNED && off < 64) r |= b & ~ones64(off);
    return r & word;
}

/* ------------------------------------------------------------------ */
/* Array loops: out[i] = op(a[i], b[i]) -- tests/pint_bench.cpp:90-94 loop
 * shape over SoA arrays. */

#define BIN_LOOP(T, OP)                                                          \
    do {                                                                         \
        const T* pa = (const T*)a; const T* pb = (const T*)b; T* po = (T*)out;   \
        for (size_t i = 0; i < n; ++i) po[i] = (T)binary1((OP), &m, pa[i], pb[i]); \
    } while (0)

#define BIN_OPS(T)                                                               \
    switch (op) {                                                                \
    case 0: BIN_LOOP(T, 0); break;   case 1: BIN_LOOP(T, 1); break;              \
    case 2: BIN_LOOP(T, 2); break;   case 3: BIN_LOOP(T, 3); break;              \
    case 4: BIN_LOOP(T, 4); break;   case 5: BIN_LOOP(T, 5); break;              \
    case 6: BIN_LOOP(T, 6); break;   case 7: BIN_LOOP(T, 7); break;              \
    case 8: BIN_LOOP(T, 8); break;   case 9: BIN_LOOP(T, 9); break;              \
    case 10: BIN_LOOP(T, 10); break; case 11: BIN_LOOP(T, 11); break;            \
    case 12: BIN_LOOP(T, 12); break;                                             \
    }

int pint_oracle_binary(int op, const pint_oracle_layout* l,
                       const void* a, const void* b, void* out, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (op < 0 || op >= PINT_O_NUM_BINARY) return -10;
    if (n && (!a || !b || !out)) return -11;
    switch (l->elem_bytes) {
    case 1: BIN_OPS(uint8_t); break;
    case 2: BIN_OPS(uint16_t); break;
    case 4: BIN_OPS(uint32_t); break;
    default: BIN_OPS(uint64_t); break;
    }
    return 0;
}

#define UN_LOOP(T, EXPR)                                                         \
    do {                                                                         \
        const T* pi = (const T*)in; T* po = (T*)out;                             \
        for (size_t i = 0; i < n; ++i) { uint64_t v = pi[i]; po[i] = (T)(EXPR); } \
    } while (0)

#define UN_TYPES(EXPR)                                                           \
    switch (l->elem_bytes) {                                                     \
    case 1: UN_LOOP(uint8_t, EXPR); break;                                       \
    case 2: UN_LOOP(uint16_t, EXPR); break;                                      \
    case 4: UN_LOOP(uint32_t, EXPR); break;                                      \
    default: UN_LOOP(uint64_t, EXPR); break;                                     \
    }

int pint_oracle_shift(int left, const pint_oracle_layout* l, uint64_t amount,
                      const void* in, void* out, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (n && (!in || !out)) return -11;
    if (left) { UN_TYPES(shift_left1(&m, v, amount)); }
    else      { UN_TYPES(shift_right1(&m, v, amount)); }
    return 0;
}

int pint_oracle_slice(const pint_oracle_layout* l, uint32_t start, uint32_t end,
                      const void* in, void* out, size_t n)
{
    int rc = pint_oracle_validate(l);
    if (rc) return rc;
    if (!(start < end && end <= l->nfields)) return -12; /* pint.hpp:748 */
    if (n && (!in || !out)) return -11;
    UN_TYPES(slice1(l, start, end, v));
    return 0;
}

static inline uint64_t load_elem(const void* p, uint32_t eb, size_t i)
{
    switch (eb) {
    case 1: return ((const uint8_t*)p)[i];
    case 2: return ((const uint16_t*)p)[i];
    case 4: return ((const uint32_t*)p)[i];
    default: return ((const uint64_t*)p)[i];
    }
}

static inline void store_elem(void* p, uint32_t eb, size_t i, uint64_t v)
{
    switch (eb) {
    case 1: ((uint8_t*)p)[i] = (uint8_t)v; break;
    case 2: ((uint16_t*)p)[i] = (uint16_t)v; break;
    case 4: ((uint32_t*)p)[i] = (uint32_t)v; break;
    default: ((uint64_t*)p)[i] = v; break;
    }
}

int pint_oracle_unpack(const pint_oracle_layout* l, int is_signed,
                       const void* packed, void* const* planes, int plane_bytes, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (plane_bytes != 1 && plane_bytes != 2 && plane_bytes != 4 && plane_bytes != 8) return -13;
    if (n && (!packed || !planes)) return -11;
    for (uint32_t f = 0; f < l->nfields; ++f) {
        if (!planes[f]) continue;
        for (size_t i = 0; i < n; ++i) {
            uint64_t v = load_elem(packed, l->elem_bytes, i);
            uint64_t x = is_signed ? (uint64_t)get_signed1(l, &m, f, v) : get1(l, &m, f, v);
            /* narrowed or widened to the plane type as a C cast does (get returns Integer,
             * get_signed make_signed<Integer>: pint.hpp:800, 810-811) */
            if (plane_bytes == 1) ((uint8_t*)planes[f])[i] = (uint8_t)x;
            else if (plane_bytes == 2) ((uint16_t*)planes[f])[i] = (uint16_t)x;
            else if (plane_bytes == 4) ((uint32_t*)planes[f])[i] = (uint32_t)x;
            else ((uint64_t*)planes[f])[i] = x;
        }
    }
    return 0;
}

int pint_oracle_repack(const pint_oracle_layout* l, const void* const* planes,
                       int plane_bytes, void* packed, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (plane_bytes != 1 && plane_bytes != 2 && plane_bytes != 4 && plane_bytes != 8) return -13;
    if (n && (!packed || !planes)) return -11;
    for (uint32_t f = 0; f < l->nfields; ++f)
        if (!planes[f]) return -11;
    for (size_t i = 0; i < n; ++i) {
        uint64_t vals[64];
        for (uint32_t f = 0; f < l->nfields; ++f) {
            /* plane values are signed integers converted to Integer (the field
             * constructor's parameters are value_type, pint.hpp:771-772) */
            vals[f] = plane_bytes == 1 ? (uint64_t)(int64_t)((const int8_t*)planes[f])[i]
                    : plane_bytes == 2 ? (uint64_t)(int64_t)((const int16_t*)planes[f])[i]
                    : plane_bytes == 4 ? (uint64_t)(int64_t)((const int32_t*)planes[f])[i]
                                       : ((const uint64_t*)planes[f])[i];
        }
        store_elem(packed, l->elem_bytes, i, pint_oracle_make_truncate1(l, &m, vals));
    }
    return 0;
}

/* ------------------------------------------------------------------ */

uint64_t pint_oracle_splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

int pint_oracle_fill(void* out, uint32_t elem_bytes, size_t n, uint64_t seed, uint64_t start)
{
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return -2;
    for (size_t i = 0; i < n; ++i)
        store_elem(out, elem_bytes, i, pint_oracle_splitmix64(seed + start + (uint64_t)i));
    return 0;
}

/* (sum, xor) of OP(a[i], b[i]) over i in [start, start + n) for the synthetic inputs
 * a[i] = (Integer) splitmix64(seed_a + start + i), b likewise: whole-array parity without
 * materialising the arrays (what pint_cuda_checksum gives for the device output range). */
int pint_oracle_stream_checksum(int op, const pint_oracle_layout* l, uint64_t seed_a, uint64_t seed_b,
                                uint64_t start, size_t n, uint64_t* sum, uint64_t* xr)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (op < 0 || op >= PINT_O_NUM_BINARY) return -10;
    uint64_t s = 0, r = 0;
    for (size_t i = 0; i < n; ++i) {
        const uint64_t a = pint_oracle_splitmix64(seed_a + start + (uint64_t)i) & m.word;
        const uint64_t b = pint_oracle_splitmix64(seed_b + start + (uint64_t)i) & m.word;
        const uint64_t v = binary1(op, &m, a, b);
        s += v; r ^= v;
    }
    if (sum) *sum = s;
    if (xr) *xr = r;
    return 0;
}

int pint_oracle_checksum(const void* x, uint32_t elem_bytes, size_t n, uint64_t* sum, uint64_t* xr)
{
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return -2;
    uint64_t s = 0, r = 0;
    for (size_t i = 0; i < n; ++i) {
        uint64_t v = load_elem(x, elem_bytes, i);
        s += v; r ^= v;
    }
    if (sum) *sum = s;
    if (xr) *xr = r;
    return 0;
}

int pint_oracle_pairs_reduce(int op, const pint_oracle_layout* l, const void* pairs, size_t n, uint64_t* sum)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (op < 0 || op >= PINT_O_NUM_BINARY) return -10;
    uint64_t s = 0;
    for (size_t i = 0; i < n; ++i) {
        uint64_t a = load_elem(pairs, l->elem_bytes, 2 * i);
        uint64_t b = load_elem(pairs, l->elem_bytes, 2 * i + 1);
        s += binary1(op, &m, a, b);
    }
    if (sum) *sum = s;
    return 0;
}

/* ---------------------------------------------------------------------------
 * Dense bit-stream storage.  NOT a reference format: the
 * reference always pads an element to find_appropriate_int (pint.hpp:724-734).
 * Spec (ours): element i of a layout with S = sum(Bits) occupies bits
 * [i*S, (i+1)*S) of a little-endian bit stream (bit k of the stream is bit
 * k % 8 of byte k / 8); the buffer is ceil(n*S / 128) * 16 bytes and every bit
 * at or above n*S is zero on output.  The per-element SEMANTICS stay the
 * reference's: a dense op is pack(op(unpack(a), unpack(b))) with `op` the
 * restatement above (pinned against the reference), so only the bit layout is
 * new.  tests/test_oracle_golden.py checks pack/unpack against an independent
 * big-integer model of this spec. */

size_t pint_oracle_dense_bytes(uint32_t total_bits, size_t n)
{
    /* n * S can exceed 64 bits only for absurd n; sizes here are test sizes */
    const unsigned __int128 bits = (unsigned __int128)n * total_bits;
    return (size_t)((bits + 127) / 128) * 16;
}

static uint64_t dense_get(const uint64_t* w, size_t nwords, uint32_t S, size_t i)
{
    const unsigned __int128 p = (unsigned __int128)i * S;
    const size_t k = (size_t)(p >> 6);
    const uint32_t sh = (uint32_t)(p & 63);
    uint64_t x = w[k] >> sh;
    if (sh && sh + S > 64 && k + 1 < nwords) x |= w[k + 1] << (64 - sh);
    return S >= 64 ? x : (x & ((UINT64_C(1) << S) - 1));
}

static void dense_put(uint64_t* w, size_t nwords, uint32_t S, size_t i, uint64_t v)
{
    const unsigned __int128 p = (unsigned __int128)i * S;
    const size_t k = (size_t)(p >> 6);
    const uint32_t sh = (uint32_t)(p & 63);
    if (S < 64) v &= (UINT64_C(1) << S) - 1;
    w[k] |= v << sh;
    if (sh && sh + S > 64 && k + 1 < nwords) w[k + 1] |= v >> (64 - sh);
}

int pint_oracle_dense_pack(const pint_oracle_layout* l, const void* padded, void* dense, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    const size_t nwords = pint_oracle_dense_bytes(m.total, n) / 8;
    uint64_t* w = (uint64_t*)dense;
    memset(w, 0, nwords * 8);
    for (size_t i = 0; i < n; ++i) dense_put(w, nwords, m.total, i, load_elem(padded, l->elem_bytes, i));
    return 0;
}

int pint_oracle_dense_unpack(const pint_oracle_layout* l, const void* dense, void* padded, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    const size_t nwords = pint_oracle_dense_bytes(m.total, n) / 8;
    for (size_t i = 0; i < n; ++i)
        store_elem(padded, l->elem_bytes, i, dense_get((const uint64_t*)dense, nwords, m.total, i));
    return 0;
}

int pint_oracle_dense_binary(int op, const pint_oracle_layout* l, const void* a, const void* b, void* out, size_t n)
{
    pint_oracle_masks m;
    int rc = pint_oracle_make_masks(l, &m);
    if (rc) return rc;
    if (op < 0 || op >= PINT_O_NUM_BINARY) return -10;
    const size_t nwords = pint_oracle_dense_bytes(m.total, n) / 8;
    uint64_t* w = (uint64_t*)out;
    /* out may be a or b: element i of the result is written after element i of both inputs was
     * read, and it only touches bits [i*S, (i+1)*S) -- but dense_put ORs, so clear as we go */
    for (size_t i = 0; i < n; ++i) {
        const uint64_t x = dense_get((const uint64_t*)a, nwords, m.total, i);
        const uint64_t y = dense_get((const uint64_t*)b, nwords, m.total, i);
        const uint64_t r = binary1(op, &m, x, y);
        /* clear the destination bits, then deposit */
        const unsigned __int128 p = (unsigned __int128)i * m.total;
        const size_t k = (size_t)(p >> 6);
        const uint32_t sh = (uint32_t)(p & 63);
        const uint64_t ones = m.total >= 64 ? ~UINT64_C(0) : ((UINT64_C(1) << m.total) - 1);
        w[k] &= ~(ones << sh);
        if (sh && sh + m.total > 64 && k + 1 < nwords) w[k + 1] &= ~(ones >> (64 - sh));
        dense_put(w, nwords, m.total, i, r);
    }
    /* zero the tail bits */
    {
        const unsigned __int128 end = (unsigned __int128)n * m.total;
        size_t k = (size_t)(end >> 6);
        const uint32_t sh = (uint32_t)(end & 63);
        if (k < nwords) {
            if (sh) { w[k] &= (UINT64_C(1) << sh) - 1; ++k; }
            for (; k < nwords; ++k) w[k] = 0;
        }
    }
    return 0;
}
