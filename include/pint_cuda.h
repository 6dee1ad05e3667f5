/*
 * pint_cuda.h -- C ABI of libpint_cuda.so: the batched (array) form of pint's
 * per-element operations, executed by hand-written sm_100a CUDA kernels.
 *
 * The reference (bad-ed/pint) is a header-only C++ template library with no
 * FFI boundary of its own: its whole interface is the free-function templates
 * of include/pint/pint.hpp ("pint.hpp:N" below), each one value in / one value
 * out, and its only array workload is the element loop of
 * tests/pint_bench.cpp:90-94.  Each entry point below is the array form of one
 * of those templates and says which; include/pint/cuda.hpp is the C++ template
 * layer (same names as the scalar API) that a pint user calls and that binds
 * to these symbols.  INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - A layout (a C++ type in the reference) crosses the ABI as a descriptor.
 *   - All array pointers are DEVICE pointers owned by the caller; element type
 *     is the layout's Integer (uint8/16/32/64_t); packed_int<Integer,...> is
 *     trivially copyable and the same size as Integer (pint.hpp:758-787), so
 *     arrays of either are interchangeable.  16-byte alignment is the fast
 *     path; any Integer-aligned pointer is accepted.
 *   - `out` may alias an input exactly (same address), not partially.
 *   - Calls are asynchronous on `stream` (a cudaStream_t; NULL = default
 *     stream) of the CURRENT device, never allocate, never synchronise, and
 *     are thread-safe for distinct streams.
 *   - Return value: 0 success; < 0 argument/layout error (PINT_E_*), the
 *     conditions the reference rejects at compile time; > 0 a cudaError_t.
 *   - There is no CPU fallback: without a CUDA device every compute entry
 *     point returns the CUDA error.
 */
#ifndef PINT_CUDA_H
#define PINT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PINT_CUDA_ABI_VERSION 1
#define PINT_MAX_FIELDS 64

/* packed_int<Integer, Bits0, Bits...> (pint.hpp:758-787) as data. */
typedef struct pint_layout {
    uint32_t elem_bytes;            /* sizeof(Integer): 1, 2, 4 or 8 (pint.hpp:731-734) */
    uint32_t nfields;               /* sizeof...(Bits) + 1, 1..64 */
    uint8_t  bits[PINT_MAX_FIELDS]; /* Bits0, Bits...; field 0 is least significant */
} pint_layout;

typedef void* pint_stream_t; /* cudaStream_t */

enum pint_op {
    PINT_ADD_WRAP = 0,               /* pint::add_wrap               pint.hpp:826-838  */
    PINT_ADD_UNSIGNED_SATURATE = 1,  /* pint::add_unsigned_saturate  pint.hpp:840-855  */
    PINT_ADD_SIGNED_SATURATE = 2,    /* pint::add_signed_saturate    pint.hpp:857-866  */
    PINT_SUB_WRAP = 3,               /* pint::sub_wrap               pint.hpp:870-884  */
    PINT_SUB_UNSIGNED_SATURATE = 4,  /* pint::sub_unsigned_saturate  pint.hpp:886-908  */
    PINT_SUB_SIGNED_SATURATE = 5,    /* pint::sub_signed_saturate    pint.hpp:910-919  */
    PINT_OR = 6,                     /* packed_int::operator|        pint.hpp:781      */
    PINT_AND = 7,                    /* packed_int::operator&        pint.hpp:782      */
    PINT_XOR = 8,                    /* packed_int::operator^        pint.hpp:783      */
    PINT_MIN_UNSIGNED = 9,           /* pint::min_unsigned           pint.hpp:936-950  */
    PINT_MAX_UNSIGNED = 10,          /* pint::max_unsigned           pint.hpp:952-966  */
    PINT_MIN_SIGNED = 11,            /* pint::min_signed             pint.hpp:968-985  */
    PINT_MAX_SIGNED = 12,            /* pint::max_signed             pint.hpp:987-1004 */
    PINT_NUM_BINARY_OPS = 13
};

enum pint_error {
    PINT_OK = 0,
    PINT_E_NULL = -1,          /* null layout or null array with n > 0 */
    PINT_E_ELEM_BYTES = -2,    /* Integer is not uint8/16/32/64_t (pint.hpp:727-734, 761-762) */
    PINT_E_NFIELDS = -3,       /* 0 or more than 64 fields */
    PINT_E_ZERO_WIDTH = -4,    /* a field of 0 bits */
    PINT_E_TOO_WIDE = -5,      /* sum(Bits) > 8 * sizeof(Integer) (pint.hpp:763-764) */
    PINT_E_OP = -10,           /* unknown operation code */
    PINT_E_INDEX = -12,        /* field index / slice bounds (pint.hpp:748, 802, 813) */
    PINT_E_PLANE_BYTES = -13,  /* plane element size is not 1, 2, 4 or 8 */
    PINT_E_ALIAS = -14,        /* partially overlapping in/out */
    PINT_E_ARG = -15,          /* other invalid argument */
    PINT_E_JIT = -20,          /* option "jit" = 2 and the layout could not be specialised (pint_cuda_jit_last_error) */
    PINT_E_TOO_LARGE = -21     /* more elements than a fixed-scratch entry point takes (PINT_STATIC_MAX_PAIRS) */
};

/* ---- library ---- */
int pint_cuda_abi_version(void);
const char* pint_cuda_error_string(int rc);
/* kernels launched by this library since it was loaded (all threads) */
unsigned long long pint_cuda_launch_count(void);
/* Tuning/testing knobs: "force_generic" (1 = always use the runtime-mask
 * kernels even for layouts that have compiled-in constants),
 * "ctas_per_sm" (0 = one CTA per tile, the default; k > 0 = persistent grid of SMs * k CTAs),
 * "ordered_ctas_per_sm" (k > 0: bandwidth-bound ops on large arrays run a persistent grid of
 * SMs * k CTAs that draw tiles in address order; default 0 = always one CTA per tile),
 * "host_chunk_bytes" (4 KiB .. 1 GiB) / "host_slots" (1 .. 8) (staging ring of the host-buffer path),
 * "dense_threads" / "dense_tile_bytes" (CTA size and per-operand tile of the dense bit-stream
 * kernels; 0 = chosen from the layout), "vec256" (32-byte global accesses in the stream kernels:
 * 0 never, 1 = default, where measured faster, 2 wherever the operands are 32-byte congruent), "dense_windows" (1 = layouts without compiled-in dense
 * kernels always use the element-by-element window kernels; testing), "dense_shift_stream" (1 = default: dense shifts
 * by less than 32 bits, in place or not, run as a masked funnel shift over the word stream; 0 = the whole-chunk kernels), "unpack_bulk" (unpack to int32 planes:
 * 1 = default, plane tiles staged in shared memory and written by bulk (TMA) stores; 0 = straight from
 * registers), "jit" (0 = default, 1, 2: run-time specialisation of unregistered layouts, see
 * pint_cuda_jit_compile), "pdl" (1 = default: the streaming kernels (binary, multi, fused, shift, slice, scalar, pairs, unpack, repack, the pair reduction, operator== over arrays, dense binary / pair / shift / conversions, and the run-time-specialised ones) are launched with
 * programmatic stream serialization, so back-to-back passes overlap one pass's tail with the next one's
 * launch: +4 ... 14 % on arrays up to 2^26 elements, neutral above; results and ordering are unchanged, every
 * kernel waits for the previous launch of its stream before touching memory; 0 = plain launches).
 * Returns 0 or PINT_E_ARG.
 * Options are process-wide; each one is read and written atomically, so setting one while other threads
 * launch is safe (a launch sees the old or the new value, never a torn one), but a change is not
 * synchronised with launches already being prepared on other threads. */
int pint_cuda_set_option(const char* key, long value);
long pint_cuda_get_option(const char* key);

/* ---- layouts (compile-time machinery of pint.hpp:27-373, 707-754 as functions) ---- */
int pint_cuda_layout_validate(const pint_layout* layout);
/* make_packed_int's storage choice (pint.hpp:789-795); 0 if unsupported */
uint32_t pint_cuda_storage_bytes(uint32_t total_bits);
/* mask_hiorder / mask_loorder / all_ones<sum> / detect_saturation_mask_type
 * (pint.hpp:324-348, 443-456) of one element, as the mask generator computes them */
int pint_cuda_layout_masks(const pint_layout* layout, uint64_t* H, uint64_t* L, uint64_t* A, uint32_t* sat_type);
/* 1 if kernels with compiled-in mask constants were built into the library for this layout, 2 if
 * they have been specialised at run time on the current device (option "jit"), else 0 */
int pint_cuda_layout_is_specialized(const pint_layout* layout);
/* Run-time specialisation (option "jit", default 0 = off).  The reference gives EVERY layout its mask
 * constants at compile time (pint.hpp:324-337, 443-456); the library has them built in for a handful
 * (registry_layouts.def) and runs the others with masks in the constant bank.  With "jit" = 1 or 2 the
 * first packed -> packed call (binary, multi, fused, binary_scalar, binary_pairs; or a dense bit-stream binary /
 * multi call, for layouts whose chunk is at most 19 words) on such a layout compiles the same kernel templates for
 * static_layout<Integer, Bits...> with NVRTC (libnvrtc.so.12, loaded on demand; headers embedded in
 * the library), caches the cubin on disk ($PINT_CUDA_JIT_CACHE, else ~/.cache/pint_cuda; written by rename, each entry with the size and
 * hash of its image, so a damaged or foreign file is recompiled, never loaded) and uses it
 * from then on: the call that compiles blocks for about a second.  pint_cuda_jit_compile does that
 * ahead of time for all op families of a layout on the current device, so that later calls never block.
 * "jit" = 1: a layout that cannot be specialised keeps running the run-time-mask kernels (identical
 * results); "jit" = 2: such a call returns PINT_E_JIT.  pint_cuda_jit_last_error: why (thread-local copy). */
int pint_cuda_jit_compile(const pint_layout* layout);
/* The same compilation for compute capability sm_major.sm_minor WITHOUT a device: fills the disk
 * cache only (a build machine can warm it for the machines that run). */
int pint_cuda_jit_prebuild(const pint_layout* layout, int sm_major, int sm_minor);
const char* pint_cuda_jit_last_error(void);
/* result layout of slice<start, end> (sliced_int, pint.hpp:746-754) */
int pint_cuda_slice_layout(const pint_layout* layout, uint32_t start, uint32_t end, pint_layout* result);

/* ---- packed -> packed (3 * sizeof(Integer) bytes per element) ---- */
/* out[i] = OP(a[i], b[i]); array form of the pint_op templates above. */
int pint_cuda_binary(int op, const pint_layout* layout, const void* a, const void* b, void* out,
                     size_t n, pint_stream_t stream);
/* outs[j][i] = ops[j](a[i], b[i]) for nops (1..PINT_MAX_MULTI_OPS) operations on the SAME
 * operands.  Op pairs that have a one-pass kernel (add_unsigned+add_signed saturate,
 * sub_unsigned+sub_signed saturate, add_signed+sub_signed saturate, add_unsigned+sub_unsigned
 * saturate, add_wrap+sub_wrap, min+max unsigned, min+max signed; either order) read a and b
 * once and write both results: 4 * sizeof(Integer) bytes per element instead of 6.  Other
 * ops of the list run as single passes; results are identical either way.  `outs` is a HOST
 * array of device pointers; outputs must not overlap the inputs or each other. */
#define PINT_MAX_MULTI_OPS 4
int pint_cuda_binary_multi(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                           void* const* outs, size_t n, pint_stream_t stream);
/* Same function computed as unpack -> per-field arithmetic -> repack inside one
 * kernel (planes in registers); bit-identical to pint_cuda_binary. */
int pint_cuda_fused_unpack_arith_repack(int op, const pint_layout* layout, const void* a, const void* b,
                                        void* out, size_t n, pint_stream_t stream);
/* out[i] = shift_left(in[i], amount) (pint.hpp:1006-1017) or
 * shift_right_unsigned (pint.hpp:1019-1029) when left == 0; uniform amount. */
int pint_cuda_shift(int left, const pint_layout* layout, uint64_t amount, const void* in, void* out,
                    size_t n, pint_stream_t stream);
/* out[i] = slice<start, end>(in[i]).value() (pint.hpp:923-932); same Integer. */
int pint_cuda_slice(const pint_layout* layout, uint32_t start, uint32_t end, const void* in, void* out,
                    size_t n, pint_stream_t stream);

/* ---- packed <-> planes ---- */
/* planes[f][i] = get<f>(packed[i]) (pint.hpp:799-807) or get_signed<f>
 * (pint.hpp:809-822) when is_signed.  `planes` is a HOST array of nfields
 * DEVICE pointers; a null entry skips that field.  plane_bytes is 1, 2, 4
 * (int32_t/uint32_t planes) or 8; the value is narrowed or widened to the
 * plane type as a C cast would -- with plane_bytes == sizeof(Integer) the planes
 * have exactly the reference's return types (Integer / make_signed<Integer>),
 * and e.g. RGB565 unpacks to three uint8_t channels at 2 + 3 bytes per element. */
int pint_cuda_unpack(const pint_layout* layout, int is_signed, const void* packed, void* const* planes,
                     int plane_bytes, size_t n, pint_stream_t stream);
/* packed[i] = packed_int(planes[0][i], planes[1][i], ...) -- the truncating
 * field constructor = make_truncate (pint.hpp:770-774, 592-601, 390-407).
 * Planes of 1, 2 or 4 bytes are read as signed integers and converted to the
 * Integer (as the constructor's value_type parameters are), then truncated to
 * the field; 8-byte planes are taken as they are. */
int pint_cuda_repack(const pint_layout* layout, const void* const* planes, int plane_bytes, void* packed,
                     size_t n, pint_stream_t stream);

/* ---- dense bit-stream storage (a storage format of THIS library) ----
 * The reference always pads an element to find_appropriate_int (pint.hpp:724-734).  A dense
 * stream stores element i in bits [i*S, (i+1)*S), S = sum(Bits), of a little-endian bit stream
 * (bit k of the stream is bit k % 8 of byte k / 8): S/8 bytes per element instead of
 * sizeof(Integer), elements and fields straddle word boundaries.  Per-element semantics are
 * the reference's: a dense op equals pack(OP(unpack(a), unpack(b))).
 *   - Dense buffers are DEVICE pointers, 16-byte aligned, pint_cuda_dense_bytes(layout, n)
 *     bytes long (= ceil(n*S/128)*16).  Bits at or above n*S are written as zero and are
 *     ignored on input.
 *   - layout->elem_bytes describes the PADDED side of pack/unpack only. */
size_t pint_cuda_dense_bytes(const pint_layout* layout, size_t n);
/* out = OP(a, b) on dense streams (3*S/8 bytes per element); out may be a or b exactly. */
int pint_cuda_dense_binary(int op, const pint_layout* layout, const void* a, const void* b, void* out, size_t n,
                           pint_stream_t stream);
/* outs[j] = ops[j](a, b) on dense streams, as pint_cuda_binary_multi: op pairs that have a one-pass
 * kernel read a and b once and write both results (4*S/8 bytes per element instead of 6*S/8);
 * outputs must not overlap the inputs or each other. */
int pint_cuda_dense_binary_multi(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                                 void* const* outs, size_t n, pint_stream_t stream);
/* out = shift_left(in, amount) (pint.hpp:1006-1017) or shift_right_unsigned (pint.hpp:1019-1029) when
 * left == 0, on a dense stream (2*S/8 bytes per element); uniform amount; out may be in exactly. */
int pint_cuda_dense_shift(int left, const pint_layout* layout, uint64_t amount, const void* in, void* out, size_t n,
                          pint_stream_t stream);
/* dense[i] = low S bits of padded[i]  /  padded[i] = dense[i] zero-extended to Integer. */
int pint_cuda_dense_pack(const pint_layout* layout, const void* padded, void* dense, size_t n, pint_stream_t stream);
int pint_cuda_dense_unpack(const pint_layout* layout, const void* dense, void* padded, size_t n, pint_stream_t stream);
/* get / get_signed (pint.hpp:799-822) of every field straight from a dense stream, and the
 * truncating field constructor (pint.hpp:770-774) straight into one; planes as in pint_cuda_unpack
 * (1, 2, 4 or 8 bytes: dense RGB24 <-> three uint8_t planes is interleaved <-> planar RGB). */
int pint_cuda_dense_to_planes(const pint_layout* layout, int is_signed, const void* dense, void* const* planes,
                              int plane_bytes, size_t n, pint_stream_t stream);
int pint_cuda_dense_from_planes(const pint_layout* layout, const void* const* planes, int plane_bytes, void* dense,
                                size_t n, pint_stream_t stream);

/* Multi-GPU for dense streams: contiguous shards whose starts are multiples of 128 ELEMENTS, so
 * that every shard's stream starts on a 16-byte boundary of the global stream (byte offset
 * begin[k] * S / 8) and is a valid dense stream of count[k] elements on its own; remainder to the
 * last shard.  No collective. */
int pint_cuda_dense_plan_shards(size_t n, int nshards, size_t* begin, size_t* count);
int pint_cuda_sharded_dense_binary(int op, const pint_layout* layout, int ngpu, const int* devices,
                                   const void* const* a, const void* const* b, void* const* out,
                                   const size_t* counts, const pint_stream_t* streams);

/* ---- the benchmark's exact loop shape (tests/pint_bench.cpp:90-94) ---- */
/* *sum_dev (device uint64_t) = sum_i OP(pairs[2i], pairs[2i+1]) mod 2^64 over an
 * AoS array of (first, second) Integer pairs; the benchmark's uint32_t `sum`
 * is the low 32 bits. */
int pint_cuda_pairs_reduce(int op, const pint_layout* layout, const void* pairs, size_t n,
                           uint64_t* sum_dev, pint_stream_t stream);
/* The same reduction for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Two passes -- one partial sum per CTA, then one CTA adds the partials into *sum_dev --
 * with the partials in a file-scope static device array of the library (64 KiB, held by every GPU that loads the
 * library until the process exits).  The call itself performs no cudaMalloc / cudaFree and calls no allocating
 * routine; `pairs` and `sum_dev` are still the caller's.  n > PINT_STATIC_MAX_PAIRS returns PINT_E_TOO_LARGE
 * (use pint_cuda_pairs_reduce).  Every call on a device shares the one scratch array, which keeps the previous
 * call's contents (nothing in it needs to start at zero: each call writes the slots it reads): calls in one
 * stream are ordered, calls that OVERLAP on the GPU (different streams or threads) race and must be serialised
 * by the caller.  Result identical to pint_cuda_pairs_reduce (integer sum, order-independent). */
#define PINT_STATIC_MAX_PAIRS (8192ull * 16384ull) /* 2^27 pairs: 8192 partial slots x 16384 pairs per CTA */
int pint_cuda_pairs_reduce_static(int op, const pint_layout* layout, const void* pairs, size_t n,
                                  uint64_t* sum_dev, pint_stream_t stream);
/* out[i] = OP(pairs[2i], pairs[2i+1]): the same interleaved operands (std::pair<Integer, Integer>
 * array, tests/pint_bench.cpp:22), the results kept as a plain array instead of summed.
 * 3 * sizeof(Integer) bytes per element; `out` must not overlap `pairs`. */
int pint_cuda_binary_pairs(int op, const pint_layout* layout, const void* pairs, void* out, size_t n,
                           pint_stream_t stream);

/* ---- packed_int::operator== / operator!= (pint.hpp:778-779) over arrays ---- */
/* equal_out[i] = (a[i] == b[i]) ? 1 : 0 (device bytes; may be NULL) and *not_equal_count_dev
 * (device uint64_t; may be NULL) = number of i with a[i] != b[i]; at least one of the two.  The
 * whole Integer is compared, unused high bits included, as the reference does.
 * 2 * sizeof(Integer) (+ 1) bytes per element. */
int pint_cuda_equal(const pint_layout* layout, const void* a, const void* b, size_t n, uint8_t* equal_out,
                    uint64_t* not_equal_count_dev, pint_stream_t stream);

/* ---- one operand the same for every element (2 * sizeof(Integer) bytes per element) ---- */
/* out[i] = OP(a[i], packed_int(scalar)), or OP(packed_int(scalar), a[i]) when scalar_first != 0
 * (the order matters for the sub_* ops and for the unused high bits of min / max, which come from
 * the second operand).  `scalar` is a packed_int::value() (pint.hpp:768, 776) and must fit the
 * Integer (else PINT_E_ARG); `out` may be `a`. */
int pint_cuda_binary_scalar(int op, const pint_layout* layout, const void* a, uint64_t scalar, int scalar_first,
                            void* out, size_t n, pint_stream_t stream);

/* ---- synthetic inputs and checksums ---- */
/* out[i] = (Integer) splitmix64(seed + start + i) */
int pint_cuda_fill_splitmix(void* out, uint32_t elem_bytes, size_t n, uint64_t seed, uint64_t start,
                            pint_stream_t stream);
/* result_dev[0] = sum, result_dev[1] = xor of the elements zero-extended to 64 bit */
int pint_cuda_checksum(const void* x, uint32_t elem_bytes, size_t n, uint64_t* result_dev,
                       pint_stream_t stream);

/* ---- multi-GPU partitioner: independent contiguous shards, no collective ---- */
/* Splits n elements into nshards contiguous ranges whose starts are multiples
 * of 128 bytes; the remainder goes to the last shard. */
int pint_cuda_plan_shards(size_t n, uint32_t elem_bytes, int nshards, size_t* begin, size_t* count);
/* Launches pint_cuda_binary on every device's shard (a[k], b[k], out[k] are
 * device pointers on devices[k], counts[k] elements) on streams[k] (may be
 * NULL = per-device default streams); returns after all launches, does not
 * synchronise.  Restores the caller's current device. */
int pint_cuda_sharded_binary(int op, const pint_layout* layout, int ngpu, const int* devices,
                             const void* const* a, const void* const* b, void* const* out,
                             const size_t* counts, const pint_stream_t* streams);

/* Same for the fused unpack -> arithmetic -> repack kernel (config C4: one launch per device). */
int pint_cuda_sharded_fused_unpack_arith_repack(int op, const pint_layout* layout, int ngpu, const int* devices,
                                                const void* const* a, const void* const* b, void* const* out,
                                                const size_t* counts, const pint_stream_t* streams);
/* Sharded unpack / repack: packed[k] and planes[k * nfields + f] live on devices[k];
 * plane arrays are sharded with the same element ranges as the packed array. */
int pint_cuda_sharded_unpack(const pint_layout* layout, int is_signed, int ngpu, const int* devices,
                             const void* const* packed, void* const* planes, int plane_bytes,
                             const size_t* counts, const pint_stream_t* streams);
int pint_cuda_sharded_repack(const pint_layout* layout, int ngpu, const int* devices,
                             const void* const* planes, int plane_bytes, void* const* packed,
                             const size_t* counts, const pint_stream_t* streams);

/* An array that lives on ONE device, handed to / collected from the shards of
 * pint_cuda_plan_shards(n, elem_bytes, ngpu): one cudaMemcpyPeerAsync per shard (NVLink when peer
 * access is enabled, see below), shard k on streams[k] (NULL = default streams).  Asynchronous.
 * shards[k] is a buffer of count[k] elements on devices[k]; a shard on the array's own device is
 * copied like the others (or pass the sub-range of the array itself as that shard: a copy onto
 * itself is skipped by the driver). */
int pint_cuda_scatter_shards(const void* src, int src_device, uint32_t elem_bytes, size_t n, int ngpu,
                             const int* devices, void* const* shards, const pint_stream_t* streams);
int pint_cuda_gather_shards(void* dst, int dst_device, uint32_t elem_bytes, size_t n, int ngpu,
                            const int* devices, const void* const* shards, const pint_stream_t* streams);
/* cudaDeviceEnablePeerAccess between every pair of `devices` that can reach each other (already
 * enabled is not an error).  Once per process; without it peer copies are staged by the driver. */
int pint_cuda_enable_peer_access(int ngpu, const int* devices);

/* ---- host-buffer entry points (the end-to-end path) ---- */
/* Pinned host memory for the calls below (cudaHostAlloc / cudaFreeHost). */
int pint_cuda_host_alloc(void** ptr, size_t bytes);
#define PINT_HOST_PORTABLE 1u        /* cudaHostAllocPortable */
#define PINT_HOST_WRITE_COMBINED 2u  /* cudaHostAllocWriteCombined: inputs the CPU only writes */
int pint_cuda_host_alloc_flags(void** ptr, size_t bytes, unsigned flags);
int pint_cuda_host_free(void* ptr);
/* out_host[i] = OP(a_host[i], b_host[i]) with HOST arrays: chunks are copied to
 * the current device, processed and copied back on a ring of streams so that
 * H2D, kernel and D2H overlap.  Synchronous: results are in out_host on
 * return.  Fast when the buffers are pinned (pint_cuda_host_alloc or
 * cudaHostRegister); pageable memory works at pageable-copy speed. */
int pint_cuda_binary_host(int op, const pint_layout* layout, const void* a_host, const void* b_host,
                          void* out_host, size_t n);
/* outs_host[j][i] = ops[j](a_host[i], b_host[i]) for nops (1..PINT_MAX_HOST_OPS)
 * operations on the SAME inputs: each chunk of a and b is uploaded once and every
 * op's result is downloaded, which halves the PCIe traffic per result of a
 * two-op pass such as add_unsigned_saturate + add_signed_saturate. */
#define PINT_MAX_HOST_OPS 4
int pint_cuda_binary_multi_host(int nops, const int* ops, const pint_layout* layout, const void* a_host,
                                const void* b_host, void* const* outs_host, size_t n);
/* The same over `ngpu` devices of this process: the HOST arrays (the reference's workload is a host
 * vector, tests/pint_bench.cpp:22,75) are split into the contiguous shards of pint_cuda_plan_shards,
 * shard k streams through devices[k]'s own staging ring, all rings are fed from the calling thread.
 * Every device's link and copy engines work at once; no collective, no peer traffic.  Synchronous.
 * Calls on different devices (this one or the two above) from different threads do not serialise:
 * the staging ring and its lock are per device. */
int pint_cuda_sharded_binary_multi_host(int nops, const int* ops, const pint_layout* layout, int ngpu,
                                        const int* devices, const void* a_host, const void* b_host,
                                        void* const* outs_host, size_t n);
/* Frees the staging rings of every device (slots * (2 + nops) * host_chunk_bytes of device memory
 * each, allocated on the first host-buffer call and otherwise kept for the life of the process). */
int pint_cuda_host_release(void);

#ifdef __cplusplus
}
#endif
#endif /* PINT_CUDA_H */
