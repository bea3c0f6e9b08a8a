/*
 * b2ntt.h -- C ABI of the B200 (sm_100a) additive-NTT / tower-field library (libb2ntt.so).
 *
 * This is the drop-in boundary for the hot path of shourovrm/binius-NTT.  The reference exposes
 * the path as C++ header templates plus a handful of compiled symbols (there is no FFI in the
 * reference); each entry point below names the reference interface it replaces, paths relative to
 * /root/reference/src/ulvt/.  The thin C++ shims in include/ulvt/ keep the reference's own class
 * names on top of these calls.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types cross this boundary.
 *   - every function returns 0 (B2NTT_OK) or a negative B2NTT_E* code; nothing aborts or throws.
 *   - "d_" pointers are device pointers on the plan's device, "h_" pointers are host pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream); device entry
 *     points are asynchronous on that stream.
 *   - GF(2^32) element = uint32_t; GF(2^128) element = four little-endian uint32_t limbs
 *     (README.md:43).  Bit-sliced layout = blocks of 32 consecutive elements, `width_bits` words
 *     per block, word 32*limb + b = bit b of that limb of the 32 elements (utils/bitslicing.cuh:32-47).
 *   - there is NO CPU fallback: without a CUDA device every compute entry point returns B2NTT_ECUDA.
 */
#ifndef B2NTT_H
#define B2NTT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

enum {
	B2NTT_OK = 0,
	B2NTT_EINVAL = -1,       /* bad argument (what makes AdditiveNTT::apply return false / ASSERT fire) */
	B2NTT_ENOMEM = -2,       /* host or device allocation failed */
	B2NTT_ECUDA = -3,        /* CUDA runtime error or no usable device */
	B2NTT_EUNSUPPORTED = -4, /* valid request this build does not implement */
	B2NTT_ECHECK = -5,       /* a requested self-check failed (b2ntt_sumcheck_set_recompute) */
	B2NTT_ELIMIT = -6        /* size beyond the fixed scratch of a no-allocation entry point (b2ntt_sumcheck_round_static) */
};

enum { B2NTT_LAYOUT_COMPACT = 0, B2NTT_LAYOUT_BITSLICED = 1 };

typedef struct b2ntt_plan b2ntt_plan;

/* Replaces AdditiveNTTConf<T,P>(log_h, log_rate) + AdditiveNTT<T,P>(conf)
 * (ntt/nttconf.cuh:49-60, ntt/additive_ntt.cuh:178-199): validates the configuration
 * (1 <= log_h, 0 <= log_rate <= 4, log_h + log_rate <= 32), builds the domain constants
 * (precompute_subspace_evals, additive_ntt.cuh:273-309, basis beta_i = 1 << i) and the pass schedule.
 *   field_bits  32 or 128
 *   layout      B2NTT_LAYOUT_COMPACT or B2NTT_LAYOUT_BITSLICED (input and output images)
 *   batch       number of independent columns transformed per call (>= 1); column b reads
 *               in[b * 2^log_h ...] and writes out[b * 2^(log_h+log_rate) ...]
 *   device      CUDA device ordinal */
int b2ntt_plan_create(b2ntt_plan** plan, int field_bits, int layout, int log_h, int log_rate, int batch, int device);
int b2ntt_plan_destroy(b2ntt_plan* plan);

/* Replaces the kernel launch loop of AdditiveNTT::apply (additive_ntt.cuh:222-247) on data that is
 * already in device memory.  d_in: batch * 2^log_h elements (novel-basis coefficients, natural order);
 * d_out: batch * 2^(log_h+log_rate) elements, coset-major per column: out[c * 2^log_h + i]
 * (additive_ntt.cuh:250-259).  d_out may equal d_in when log_rate == 0. */
int b2ntt_exec(b2ntt_plan* plan, const void* d_in, void* d_out, void* stream);

/* b2ntt_exec with CUDA events recorded on `stream` around every pass launch (measurement aid, no
 * reference counterpart).  Synchronises; writes the device time of each pass in milliseconds to
 * pass_ms[0..n) and, when pass_kind is not NULL, the kernel of each pass (0: b2::pass_kernel, the
 * bundle pass through shared memory; 1: b2::upper_kernel, the warp-uniform-twiddle pass; 2: b2::bottom_kernel, the 512-line tile pass).  Returns the
 * number of passes n (<= max_passes) or a negative error. */
int b2ntt_exec_profile(b2ntt_plan* plan, const void* d_in, void* d_out, void* stream, float* pass_ms, int* pass_kind, int max_passes);

/* Inverse of b2ntt_exec on ONE coset per column (no reference counterpart):
 * d_in = batch * 2^log_h evaluations on coset `coset`, d_out = batch * 2^log_h coefficients. */
int b2ntt_exec_inverse(b2ntt_plan* plan, const void* d_in, void* d_out, int coset, void* stream);

/* Forward transform of ONE coset row per column: d_in = batch * 2^log_h coefficients, d_out = batch *
 * 2^log_h evaluations on coset `coset` (0 <= coset < 2^log_rate) -- row `coset` of b2ntt_exec's output.
 * No reference counterpart (the reference always produces every coset, additive_ntt.cuh:213-215); this
 * is phase 2 of a transform split over several devices: device h of 2^g runs coset h of a
 * (log_h - g, log_rate = g) plan on the block it owns.  d_out may equal d_in. */
int b2ntt_exec_coset(b2ntt_plan* plan, const void* d_in, void* d_out, int coset, void* stream);

/* Phase 1 of a transform split over 2^log_segments devices: `plan` describes the
 * FULL transform (log_rate 0); d_buf holds this device's 2^log_segments segments of
 * 2^(log_h - 2*log_segments) elements, segment h = global indices (h, m = this device, .).  Runs the top
 * log_segments butterfly stages (additive_ntt.cuh:127-159 for stages log_h-1 .. log_h-log_segments) in
 * place; their twiddles depend only on the h bits above the butterfly bit.  1 <= log_segments <= 4. */
int b2ntt_exec_top(b2ntt_plan* plan, void* d_buf, int log_segments, void* stream);

/* Replaces bool AdditiveNTT::apply(const NTTData&, NTTData&) (additive_ntt.cuh:201-265): host input,
 * host output, H2D + transform + D2H + synchronise on the plan's own stream.  n_in is the input
 * element count per column and must equal 2^log_h (else B2NTT_EINVAL, apply()'s `return false`). */
int b2ntt_exec_host(b2ntt_plan* plan, const void* h_in, size_t n_in, void* h_out);

/* `count` independent calls of AdditiveNTT::apply in one pipelined call: the host-to-device copy of transform i,
 * the kernels of transform i-1 and the device-to-host copy of transform i-2 run on three streams over three
 * device buffer sets (PCIe is full duplex), so the host path runs at the speed of the slowest of the three
 * instead of their sum.  h_in[i] / h_out[i]: host buffers as for b2ntt_exec_host (pinned memory for full speed;
 * h_out[i] must not be read before the call returns; buffers may repeat among i only if that is harmless to the
 * caller).  Synchronises before returning.  No reference counterpart (apply() is one blocking call, :262). */
int b2ntt_exec_host_batch(b2ntt_plan* plan, const void* const* h_in, size_t n_in, void* const* h_out, int count);

/* Query: number of kernel launches one b2ntt_exec issues, and HBM passes over the data set. */
int b2ntt_plan_launches(const b2ntt_plan* plan);
int b2ntt_plan_passes(const b2ntt_plan* plan);
/* Copies the plan's normalised domain constants C[i][j] (row pitch 32 words, 32 rows) to h_out. */
int b2ntt_plan_constants(const b2ntt_plan* plan, uint32_t* h_out);

/* Replaces transpose_kernel<W> / untranspose_kernel<W> (utils/bitslicing.cuh:89-104), in place on
 * n_blocks blocks of 32 elements.  width_bits 32, 64 or 128.  dir 0: compact -> bit-sliced,
 * 1: bit-sliced -> compact. */
int b2ntt_bitslice(void* d_buf, size_t n_blocks, int width_bits, int dir, void* stream);

/* Element-wise tower product out[i] = a[i] * b[i], n elements.
 *   compact   field_bits 8/16/32 (one element per uint32_t, FanPaarTowerField<H>::multiply,
 *             finite_fields/binary_tower.cuh:113-115) or 128 (4 limbs, sumcheck/test/utils/tower_7_mul.cu:4-20)
 *   bitsliced field_bits 32 or 128, n a multiple of 32 (multiply_unrolled<5|7>,
 *             finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh:4-5)
 * d_out may alias d_a or d_b. */
int b2ntt_mul(int field_bits, int layout, const void* d_a, const void* d_b, void* d_out, size_t n, void* stream);

/* Dependent-chain bit-sliced multiply microbenchmark: every thread squares-and-multiplies `iters`
 * times in registers (the shape of bitsliced_repeat, finite_fields/tests/profiling/kernels/
 * bitsliced_repeat.cu:6-33).  Used to measure the logic-pipe peak.  d_io: threads * 2^h words.
 * field_bits 32: x <- x * y on 32-word operands in registers; 128: the same chain on 128-word blocks
 * (multiply_unrolled<7> per step), operands in thread-private rows of shared memory.  In both forms
 * y = the block's words rotated by 33 (32: by 1) and XORed with 0x9e3779b9 * (index + 1). */
int b2ntt_mul_repeat(int field_bits, void* d_io, size_t n_threads, int iters, void* stream);

/* Replaces bool check_gpu_capabilities() (utils/common.cu:6-43): 1 capable, 0 not, <0 error. */
int b2ntt_check_gpu_capabilities(int device);

/* ---- bit-sliced GF(2^128) sumcheck prover for a product of `composition_size` multilinears ----
 * Replaces Sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED> (sumcheck/sumcheck.cuh:10-300) and its kernels
 * compute_compositions (sumcheck/core/kernels.cuh:5-102) / fold_large_list_halves (sumcheck/core/kernels.cu:9-30).
 * num_vars 5..30 (the reference instantiates 20, 24, 28), composition_size 2..4. */
typedef struct b2ntt_sumcheck b2ntt_sumcheck;
int b2ntt_sumcheck_create(b2ntt_sumcheck** sc, int num_vars, int composition_size, int device);
int b2ntt_sumcheck_destroy(b2ntt_sumcheck* sc);
/* The constructor's upload (sumcheck.cuh:79-126): composition_size columns of 2^num_vars GF(2^128) evaluations, compact
 * (transposed = 0: bit-sliced on the device first, transpose_kernel<128>) or already bit-sliced (transposed = 1).
 * load_host copies into a buffer the object owns; load_device adopts the caller's device buffer, which the protocol
 * then overwrites in place.  Either call restarts the protocol at round 0. */
int b2ntt_sumcheck_load_host(b2ntt_sumcheck* sc, const void* h_evals, int transposed);
int b2ntt_sumcheck_load_device(b2ntt_sumcheck* sc, void* d_evals, int transposed);
/* this_round_messages (sumcheck.cuh:136-246): sum[4] = XOR over the remaining hypercube of the product; points[(C+1)*4] =
 * the round polynomial at t = 0..C.  Valid for rounds 0..num_vars (the last call returns the final evaluation claim). */
int b2ntt_sumcheck_round_messages(b2ntt_sumcheck* sc, uint32_t* sum, uint32_t* points);
/* move_to_next_round (sumcheck.cuh:248-299): binds the top remaining variable to challenge[4]. */
int b2ntt_sumcheck_next_round(b2ntt_sumcheck* sc, const uint32_t* challenge);
/* From the second round on p(1) is normally DERIVED (p(1) = q(r) + p(0), q = the previous round polynomial, r = the last
 * challenge) instead of recomputed from the data.  on != 0: every device round computes p(1) from the data as the reference
 * does (compute_compositions, sumcheck/core/kernels.cuh:5-102) and round_messages returns B2NTT_ECHECK if it differs from
 * the derived value -- the check a caller (or a test) switches on to make the per-round identities non-tautological. */
int b2ntt_sumcheck_set_recompute(b2ntt_sumcheck* sc, int on);
int b2ntt_sumcheck_round(const b2ntt_sumcheck* sc);
long long b2ntt_sumcheck_launches(const b2ntt_sumcheck* sc);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The messages of the FIRST round --
 * what b2ntt_sumcheck_round_messages returns right after b2ntt_sumcheck_load_device(.., transposed = 1) -- computed with no
 * object and no allocation of any kind: the accumulator of the round kernel is a static __device__ array of
 * (B2NTT_STATIC_ROUND_MAX_COMPOSITION + 1) x 128 words (2560 bytes, held by every GPU that loads the library until exit).
 *   d_evals : composition_size columns of 2^num_vars GF(2^128) evaluations, bit-sliced, read only (device memory)
 *   d_points: (composition_size + 1) x 4 words, d_sum: 4 words (device memory of the same GPU); written on `stream`,
 *             valid once the stream reaches that point -- the call does not synchronise and copies nothing to the host.
 * num_vars 6..B2NTT_STATIC_ROUND_MAX_VARS, composition_size 2..B2NTT_STATIC_ROUND_MAX_COMPOSITION; beyond the maximum:
 * B2NTT_ELIMIT, with the reason in b2ntt_last_cuda_error().  All calls on a GPU share the one array (cleared at the start
 * of each call): calls that overlap on the GPU -- two streams, two graph branches -- race; serialise them. */
#define B2NTT_STATIC_ROUND_MAX_COMPOSITION 4
#define B2NTT_STATIC_ROUND_MAX_VARS 30
int b2ntt_sumcheck_round_static(const void* d_evals, int num_vars, int composition_size, void* d_points, void* d_sum, void* stream);

/* ---- additive NTT over GF(2^128) on a caller-supplied basis (full-width twiddles / general domains) ----
 * Generalises AdditiveNTT<T,P> (ntt/additive_ntt.cuh:176-309), whose basis is fixed to beta_i = 1 << i (:281-283) and whose
 * twiddles therefore stay in GF(2^32).  h_basis = log_h elements of four little-endian limbs, linearly independent over
 * GF(2) (B2NTT_EINVAL otherwise); NULL = the reference basis, for which the output equals b2ntt_exec's (log_h <= 32).
 * Input = novel-basis coefficients in natural order, output = evaluations, point i = XOR of beta_b over the set bits b of
 * i (additive_ntt.cuh:250-259).  1 <= log_h <= 34; log_rate is not separate here: a rate-1/R code is a log_h + log R basis
 * with zero-padded input. */
typedef struct b2ntt_general b2ntt_general;
int b2ntt_general_create(b2ntt_general** g, int log_h, const uint32_t* h_basis, int device);
int b2ntt_general_destroy(b2ntt_general* g);
/* d_in / d_out: 2^log_h compact GF(2^128) elements on the device (may alias); inverse != 0 runs evaluations -> coefficients. */
int b2ntt_general_exec(b2ntt_general* g, const void* d_in, void* d_out, int inverse, void* stream);
/* The normalised domain constants W^_i(beta_{i+1+j}) (precompute_subspace_evals, additive_ntt.cuh:273-309), row-major
 * [log_h][log_h] elements of 4 words, row i holding log_h - 1 - i entries (rest zero). */
int b2ntt_general_constants(const b2ntt_general* g, uint32_t* h_out);

/* ---- multi-GPU on one NVLink / NVSwitch box (no reference counterpart: the reference is single-GPU) ----
 * ONE transform of 2^log_n elements over 2^log_ranks devices ("four-step", BASELINE config 4).  Global index
 * i = (h : g bits, m : g bits, l).  Rank m starts with the 2^g segments (h, m, .) ("m-shard") and ends with the contiguous
 * block [m * 2^(log_n-g), ...) of the transform in natural order.  The reference runs its layers from the highest stride down
 * (additive_ntt.cuh:233-247), so there is exactly one exchange, and it is FUSED into phase 1: the kernel that runs the top g
 * layers stores segment h straight into rank h's block buffer with 16-byte peer stores over NVLink (no collective call).
 * One object per rank; the ranks may live in one process (peer access) or in one process each (CUDA IPC, helpers below).
 * Caller's protocol per transform: [all blocks free of readers] -> scatter on every rank -> [every scatter complete] ->
 * local on every rank.  b2ntt_mg_exec does this with events; a torchrun caller uses two stream-ordered barriers. */
typedef struct b2ntt_fourstep b2ntt_fourstep;
int b2ntt_fourstep_create(b2ntt_fourstep** fs, int field_bits, int log_n, int log_ranks, int rank, int device);
int b2ntt_fourstep_destroy(b2ntt_fourstep* fs);
size_t b2ntt_fourstep_block_bytes(const b2ntt_fourstep* fs);   /* bytes of a rank's shard = of its block */
int b2ntt_fourstep_launches(const b2ntt_fourstep* fs);
/* d_block = rank peer_rank's block buffer as mapped in THIS process (all 2^log_ranks of them, the own one included) */
int b2ntt_fourstep_set_peer(b2ntt_fourstep* fs, int peer_rank, void* d_block);
int b2ntt_fourstep_scatter(b2ntt_fourstep* fs, const void* d_shard, void* stream);   /* phase 1 + exchange; d_shard is read only */
int b2ntt_fourstep_local(b2ntt_fourstep* fs, void* d_block, void* stream);           /* phase 2, in place on the own block */
/* CUDA IPC for one-process-per-GPU callers: handle64 = 64 bytes (cudaIpcMemHandle_t of the allocation that holds d_ptr),
 * offset = d_ptr's offset inside it; open maps it into the calling process on `device` (peer access enabled lazily). */
int b2ntt_ipc_export(const void* d_ptr, void* handle64, size_t* offset);
int b2ntt_ipc_open(const void* handle64, size_t offset, int device, void** d_ptr);
int b2ntt_ipc_close(void* d_ptr, size_t offset, int device);

/* Single-process drivers for a caller that owns all the GPUs (b2ntt_mg_create / exec / destroy).
 * B2NTT_MG_COLUMN_SHARD (config 3): `batch` independent columns of (log_h, log_rate), contiguous slabs per device
 * (b2ntt_mg_columns), no data moves between devices.  B2NTT_MG_FOUR_STEP (config 4): log_rate = 0, batch = 1, n_gpus a power
 * of two; d_in[i] = rank i's m-shard (read only), d_out[i] = block i.  exec launches on one stream per device and returns
 * when every device has finished. */
#define B2NTT_MG_COLUMN_SHARD 0
#define B2NTT_MG_FOUR_STEP 1
typedef struct b2ntt_mg b2ntt_mg;
int b2ntt_mg_create(b2ntt_mg** mg, int mode, int n_gpus, const int* devices, int field_bits, int log_h, int log_rate, int batch);
int b2ntt_mg_columns(const b2ntt_mg* mg, int i, int* first, int* count);
int b2ntt_mg_exec(b2ntt_mg* mg, void* const* d_in, void* const* d_out);
int b2ntt_mg_destroy(b2ntt_mg* mg);

/* ---- BabyBear (p = 15 * 2^27 + 1) radix-2 NTT ----
 * Replaces NTT<BB31> (ntt/gpuntt.cuh:126-209) with its configuration NTTConfRad2<BB31> (ntt/nttconf.cuh:25-47), its
 * kernels ntt_kernel (gpuntt.cuh:65-124) and bit_reverse_in_place_ker (:22-34).  Words are risc0::Fp RAW values
 * (Montgomery form, R = 2^32; finite_fields/risc0_baby_bear.h:163-177), i.e. the memory image of a BB31 array; the
 * generator is given the same way.  1 <= log_n <= 27, log_group_order >= log_n (nttconf.cuh:32-39).
 * Output = the reference's: an IN_ORDER input is bit-reversed first, then stages 0 .. log_n-1 of u + v, (u - v) * w. */
typedef struct b2ntt_bb31 b2ntt_bb31;
int b2ntt_bb31_create(b2ntt_bb31** out, uint32_t generator_raw, int log_group_order, int log_n, int device);
int b2ntt_bb31_destroy(b2ntt_bb31* h);
/* device-resident, asynchronous on `stream`; d_out may equal d_in only when the input is already bit-reversed */
int b2ntt_bb31_exec(b2ntt_bb31* h, const void* d_in, void* d_out, int input_bit_reversed, void* stream);
/* NTT<E>::apply (gpuntt.cuh:158-183): host in (n = 2^log_n elements), host out, synchronous */
int b2ntt_bb31_exec_host(b2ntt_bb31* h, const void* h_in, size_t n, int input_bit_reversed, void* h_out);
int b2ntt_bb31_launches(const b2ntt_bb31* h);

const char* b2ntt_strerror(int code);
/* Text of the last CUDA error seen by the calling thread's last failing call ("" if none). */
const char* b2ntt_last_cuda_error(void);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
