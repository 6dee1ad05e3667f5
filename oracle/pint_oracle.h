/*
 * pint_oracle.h -- CPU restatement of the reference's per-element algorithms.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.  The product path (libpint_cuda.so)
 * never links, loads or calls it and has no CPU fallback.
 *
 * Parity status: PINNED.  The port is checked against (a) every known-answer
 * test of the reference's own tests/pint_test.cpp (tests/golden/kat_pint_test.json,
 * extracted by tests/golden/gen_golden.py), (b) the README examples and
 * (c) outputs of the unmodified reference header compiled into
 * oracle/_ref/libpint_ref.so (tests/golden/ref_vectors.npz and, when the .so
 * is present, live differential runs).
 *
 * All citations "pint.hpp:N" are to /root/reference/include/pint/pint.hpp.
 *
 * The reference expresses a layout as a C++ type (packed_int<Integer, Bits...>).
 * Here a layout is a runtime descriptor; every function computes the same
 * expression the header would instantiate for that type, in uint64_t arithmetic
 * truncated to the element width (which is what Integer promotion followed by
 * the packed_int constructor's truncation amounts to, pint.hpp:547,581,589,768).
 */
#ifndef PINT_ORACLE_H
#define PINT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same shape as pint_layout in include/pint_cuda.h. */
typedef struct pint_oracle_layout {
    uint32_t elem_bytes;  /* sizeof(Integer): 1, 2, 4 or 8 (pint.hpp:731-734) */
    uint32_t nfields;     /* sizeof...(Bits) + 1, 1..64 */
    uint8_t  bits[64];    /* Bits0, Bits... ; field 0 is the least significant */
} pint_oracle_layout;

/* Operation codes -- identical numbering to include/pint_cuda.h. */
enum {
    PINT_O_ADD_WRAP = 0,
    PINT_O_ADD_UNSIGNED_SATURATE = 1,
    PINT_O_ADD_SIGNED_SATURATE = 2,
    PINT_O_SUB_WRAP = 3,
    PINT_O_SUB_UNSIGNED_SATURATE = 4,
    PINT_O_SUB_SIGNED_SATURATE = 5,
    PINT_O_OR = 6,
    PINT_O_AND = 7,
    PINT_O_XOR = 8,
    PINT_O_MIN_UNSIGNED = 9,
    PINT_O_MAX_UNSIGNED = 10,
    PINT_O_MIN_SIGNED = 11,
    PINT_O_MAX_SIGNED = 12,
    PINT_O_NUM_BINARY = 13
};

/* Derived per-layout constants (what the header's L0 metafunctions yield). */
typedef struct pint_oracle_masks {
    uint64_t word;        /* all ones over the Integer (8*elem_bytes bits) */
    uint64_t H;           /* mask_hiorder, pint.hpp:324-329 */
    uint64_t L;           /* mask_loorder, pint.hpp:332-337 */
    uint64_t A;           /* all_ones<sum(Bits)>, pint.hpp:340-348 */
    uint32_t total;       /* sum(Bits), pint.hpp:257-271 */
    uint32_t maxbits;     /* find_max, pint.hpp:277-285 */
    uint32_t sat_type;    /* detect_saturation_mask_type, pint.hpp:443-456 */
    uint32_t nuniq;       /* number of unique widths, first-appearance order (pint.hpp:97-110) */
    uint8_t  off[64];     /* mask_offsets_vector, pint.hpp:288-292 */
    uint8_t  uniq_w[64];  /* unique widths */
    uint64_t uniq_lo[64]; /* per unique width: LSB mask of the fields of that width
                             (unsigned_saturation_mask_type_2, pint.hpp:503-515) */
} pint_oracle_masks;

/* 0 on success, negative on an invalid layout (the conditions the header
 * rejects at compile time: pint.hpp:761-764, 727-728). */
int pint_oracle_validate(const pint_oracle_layout* l);
int pint_oracle_make_masks(const pint_oracle_layout* l, pint_oracle_masks* m);

/* smallest storage width in bytes for a given sum of bits (make_packed_int,
 * pint.hpp:789-795 with clp2 :710-722 and find_appropriate_int :724-734);
 * 0 when unsupported (> 64 or 0). */
uint32_t pint_oracle_storage_bytes(uint32_t total_bits);

/* ---- single element (values are carried in uint64_t, truncated to Integer) ---- */
uint64_t pint_oracle_binary1(int op, const pint_oracle_masks* m, uint64_t a, uint64_t b);
uint64_t pint_oracle_shift_left1(const pint_oracle_masks* m, uint64_t v, uint64_t amount);
uint64_t pint_oracle_shift_right_unsigned1(const pint_oracle_masks* m, uint64_t v, uint64_t amount);
uint64_t pint_oracle_get1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t index, uint64_t v);
int64_t  pint_oracle_get_signed1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t index, uint64_t v);
uint64_t pint_oracle_slice1(const pint_oracle_layout* l, const pint_oracle_masks* m, uint32_t start, uint32_t end, uint64_t v);
uint64_t pint_oracle_make_truncate1(const pint_oracle_layout* l, const pint_oracle_masks* m, const uint64_t* values);

/* ---- arrays: out[i] = f(a[i], b[i]) in the loop shape of
 * tests/pint_bench.cpp:90-94 rewritten over SoA arrays.  Element type is the
 * layout's Integer.  All return 0 or a negative argument error. ---- */
int pint_oracle_binary(int op, const pint_oracle_layout* l,
                       const void* a, const void* b, void* out, size_t n);
int pint_oracle_shift(int left, const pint_oracle_layout* l, uint64_t amount,
                      const void* in, void* out, size_t n);
int pint_oracle_slice(const pint_oracle_layout* l, uint32_t start, uint32_t end,
                      const void* in, void* out, size_t n);
/* planes[f] may be NULL (field skipped).  plane_bytes is 4 or 8; a field value
 * is narrowed to the plane type exactly as a C cast would (lossless when
 * bits[f] <= 8*plane_bytes). */
int pint_oracle_unpack(const pint_oracle_layout* l, int is_signed,
                       const void* packed, void* const* planes, int plane_bytes, size_t n);
int pint_oracle_repack(const pint_oracle_layout* l, const void* const* planes,
                       int plane_bytes, void* packed, size_t n);

/* Independent naive model of a binary op (extract each field, do the arithmetic
 * in wide integers, clamp or wrap, re-insert; bits above sum(Bits) follow the
 * reference's rule: zero for add/sub, pass-through for | & ^, taken from b for
 * min/max).  Used only to cross-check the SWAR port above. */
uint64_t pint_oracle_naive_binary1(int op, const pint_oracle_layout* l, uint64_t a, uint64_t b);

/* Synthetic inputs: x[i] = (Integer) splitmix64(seed + start + i). */
uint64_t pint_oracle_splitmix64(uint64_t x);
int pint_oracle_fill(void* out, uint32_t elem_bytes, size_t n, uint64_t seed, uint64_t start);

/* Checksum of an element array: sum and xor of the elements (zero-extended to
 * 64 bit), the same pair libpint_cuda's pint_cuda_checksum produces. */
int pint_oracle_checksum(const void* x, uint32_t elem_bytes, size_t n, uint64_t* sum, uint64_t* xr);
/* (sum, xor) of OP(a[i], b[i]), i in [start, start + n), a / b the splitmix inputs above */
int pint_oracle_stream_checksum(int op, const pint_oracle_layout* l, uint64_t seed_a, uint64_t seed_b,
                                uint64_t start, size_t n, uint64_t* sum, uint64_t* xr);

/* AoS pair reduction of tests/pint_bench.cpp:90-94: sum += op(P(first), P(second)).value()
 * over (first, second) pairs of Integer, accumulated in the Integer's width. */
int pint_oracle_pairs_reduce(int op, const pint_oracle_layout* l, const void* pairs, size_t n, uint64_t* sum);

/* ---- dense bit-stream storage (OUR format, see pint_oracle.c) ----
 * element i occupies bits [i*S, (i+1)*S) of a little-endian bit stream, S = sum(Bits);
 * buffers are ceil(n*S/128)*16 bytes, bits at or above n*S are zero on output.
 * dense_binary = pack(op(unpack(a), unpack(b))) with op the restatement above. */
size_t pint_oracle_dense_bytes(uint32_t total_bits, size_t n);
int pint_oracle_dense_pack(const pint_oracle_layout* l, const void* padded, void* dense, size_t n);
int pint_oracle_dense_unpack(const pint_oracle_layout* l, const void* dense, void* padded, size_t n);
int pint_oracle_dense_binary(int op, const pint_oracle_layout* l, const void* a, const void* b, void* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif
