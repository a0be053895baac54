/*
 * fricc_b200.h -- C-ABI of libfricc_b200.so: the B200 (sm_100a) implementation of fri-cc's FRI-CCSD inner loop.
 *
 * This is the drop-in boundary.  Every entry point below replaces one symbol of the reference's pybind11
 * module `py_rccsd` (/root/reference/src/py_rccsd.cpp:16-55) or one of the C++ functions behind it; the
 * reference interface each one stands in for is cited next to it.  Signatures use plain pointers, sizes and
 * NUL-terminated strings only -- no torch, numpy or C++ types -- so the library can be bound from ctypes
 * (fricc_b200/py_rccsd.py does), pybind11, cffi or anything else.  INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every function returns an int status: FRICC_OK (0) or a negative FRICC_E* code; fricc_last_error() returns a
 *     human-readable message for the calling thread's last failure.  Nothing here ever calls exit(): the
 *     conditions under which the reference prints to stderr and exits become error codes.
 *   - array arguments carry a `mem` tag: FRICC_HOST (ordinary host memory; the library stages it through the GPU,
 *     copies are part of the call) or FRICC_DEVICE (a CUDA device pointer on the current device; zero-copy).
 *   - all work is issued on the stream set with fricc_set_stream() (default: the legacy default stream) and every
 *     call that returns results to host memory synchronises that stream before returning.
 *   - indices are int64_t (the reference uses size_t), values are IEEE binary64.
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point returns FRICC_ECUDA.
 */
#ifndef FRICC_B200_H
#define FRICC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FRICC_OK 0
#define FRICC_EINVAL (-1)    /* bad argument (unknown term/compression/method string, m > N, ndim, null ptr) */
#define FRICC_ECUDA (-2)     /* CUDA runtime error or no device                                               */
#define FRICC_EPROB (-3)     /* a probability exceeded 1 (reference: fri_utils.cpp:383-395 exits)             */
#define FRICC_EPIVOTAL (-4)  /* pivotal acceptance probability out of [0,1] (fri_utils.cpp:84-94 exits)       */
#define FRICC_ECOUNT (-5)    /* sampler produced the wrong number of samples (fri_utils.cpp:110-124 exits)    */
#define FRICC_EDRAWS (-6)    /* caller-supplied draw buffer too short                                         */
#define FRICC_ERANGE (-7)    /* element index out of range                                                    */

#define FRICC_HOST 0
#define FRICC_DEVICE 1

typedef struct fricc_st* fricc_st_t; /* opaque SparseTensor4d handle (device-resident COO + cached bucketed layouts) */

/* ---- library / runtime ------------------------------------------------------------------------------------- */
const char* fricc_last_error(void);
const char* fricc_version(void);
int fricc_device_count(int* count);
int fricc_set_device(int device);
int fricc_set_stream(void* cuda_stream);      /* cudaStream_t of the caller (e.g. torch's current stream); NULL = default */
int fricc_stream_synchronize(void);
/* tuning / A-B switches.  "d_candidates": first size of the candidate list ordered when looking for D (default 65536;
 * it grows x16 until the greedy test stops inside the list -- any value gives identical results); "d_estimate": 1
 * (default) sizes that list from a count+sum histogram estimate of where the greedy test stops, 0 = geometric growth only.
 * "escan": 1 (default) = exact running sums by the parallel binade-segmented scan,
 * 0 = by the one-thread chain kernels (bit-identical results, kept as the in-library cross-check).
 * "push_ctas_per_sm": grid cap of the peer-store kernels (default 2 CTAs per SM: NVLink-bound, and it leaves thread
 * slots and registers to the contraction kernel that runs concurrently on the compute stream); "push_carveout": the
 * L1/shared split (percent shared) those kernels ask for, default 100 = the split the tensor-core kernel runs under,
 * so that both can be resident on one SM; -1 = the device default.  "push_copy_engine": 1 = fricc_push_view
 * moves views that collapse to contiguous rows of at least "push_min_row_bytes" (default 4096) with one
 * cudaMemcpy(2D)Async per peer (copy engines, no SM); 0 (default) = always the store kernel (same bytes in the same
 * places; measured faster for the step as a whole, see fricc_peer.cu).
 * "alloc_cache_mb": bytes of freed per-call scratch (sign matrices, CSR remainders, sort buffers; blocks up to 2 GB) the
 * library keeps per device (each block with an event recorded at its free; the stream that takes it out again waits for
 * that event, so reuse is ordered whichever stream asks) and hands out again by size instead of going back to cudaMallocAsync -- default
 * 24576, 0 = none.  A steady-state iteration then makes no allocation call at all: a cudaMallocAsync that has to map
 * memory enters the kernel driver and queues behind e.g. `nvidia-smi -lms` queries (measured: 0.03 - 2.7 s host stalls).
 * "release_cached" (any value): give every cached block back now.  Stats "alloc_driver_calls" / "alloc_cache_hits".
 * Environment: FRICC_TRACE_SLOW_MS=<ms> reports every runtime call that keeps the host longer than that on stderr.  */
int fricc_set_option(const char* key, int value);
/* Two contractions into the SAME output with the same T as ONE dense product where the algebra allows it (an extension;
 * the reference's drivers make the two calls one after the other, e.g. fricc/friccsd.py:347-356 (0112, 0113 -> Wvoov) and
 * :252-258 (1302, 1202 -> tmp)).  All arrays on the device.  Fused when the output maps agree, the prefactors are in a
 * small-integer ratio and either
 *   - both terms read the same operand view and differ in which axes of T form (R, K): the signs of both are scattered,
 *     weighted, into one int8 matrix  (0112 + 0113:  out += 0.5 (2 A_0113 - A_0112) W), or
 *   - both terms take (R, K) from the same axes of T and read different operands: planes of the combined panel
 *     (1302 + 1202:  out += A (2 Wvoov - Wvovo); 0113 + 0313 likewise);
 * otherwise -- or when the tensor-core form would not be chosen -- exactly the two fricc_contract calls.  The result
 * equals the two calls up to fp64 rounding of the combination (tests/test_plugin_tc_gpu.py).
 * fricc_get_stat("fused_pairs") counts the calls that ran fused.                                                         */
int fricc_contract_pair(const double* W_a_dev, const int64_t a_shape[4], const int64_t a_strides[4], const double* W_b_dev,
                        const int64_t b_shape[4], const int64_t b_strides[4], fricc_st_t T, double* out_dev,
                        const int64_t out_shape[4], const int64_t out_strides[4], const char* term_a, const char* term_b);
/* "spmm_exact_order": 0 (default) = the fast contraction kernels ((R,K)-sorted buckets, FMA, within 1e-10 of the
 * reference); 1 = PARITY MODE: every contraction adds the entries of an owner pair in storage order with one rounded
 * product and one rounded sum per entry straight onto the output element, exactly like the reference's loop nest
 * (sparse_tensor.cpp:66-73) -- bit-identical outputs, fp64 SpMM only (the tensor-core form is bypassed).
 * "tc_mode": which form fricc_contract picks -- 0 = always the fp64 SpMM, 1 (default) = the tensor-core form when its
 * cost model says it wins (an FRI-compressed T with at least half of its entries sampled, a large enough product),
 * 2 = always the tensor-core form.  "tc_planes": digit planes it expands W into (default 6 = 48 bits per column).
 * "tc_predict" (1: see fricc_wslices_redone), "tc_predict_persistent" (0: the re-encoding kernel runs one 64 KB unit per block so that kernels of another stream get in between; 1: persistent blocks), "tc_pair" (1: CTA-pair tensor-core kernel, 0: one CTA per tile), "tc_lockstep" (1) with "tc_sync_shift" (4: a milestone every 16 k-blocks) and "tc_sync_lag" (2), "tc_band", "tc_fused_planes": 1 (default) = operands whose contracted run is contiguous (the ladder's native W[a,b,c,d]) are
 * expanded by the single-read cluster kernel, 0 = by the two-pass kernels (identical planes).
 * "tc_plane_cache": 1 = fricc_contract keeps the planes of up to 12 device operands across calls, keyed by address,
 * shape and strides and guarded by a fingerprint of 4096 sampled elements (for drivers that hand over the same
 * operands every iteration, fricc/frilccsd.py:171-179); default 0 = planes are rebuilt on every call.               */
int fricc_plane_cache_clear(void);
/* Declare the n_elements doubles at device address p CONSTANT until further notice (on = 1) or withdraw that (on = 0).
 * The ERI blocks of a coupled-cluster driver never change between iterations -- fricc/friccsd.py:347-378 hands
 * eris.ovov to five of the eleven terms in every cycle -- so the digit planes of any view that starts inside a declared
 * range are built on first use and kept, keyed by (address, shape, strides, contracted/free axes, planes); no
 * fingerprint and no read-back, the declaration is the guarantee.  Declaring a range again (after rewriting it) or
 * withdrawing it drops its planes; so does fricc_plane_cache_clear.  Stat "constant_plane_hits".  An extension of the
 * plugin surface (py_rccsd.constant_operand); results are bit-identical to undeclared calls.                         */
int fricc_constant_operand(const double* p, int64_t n_elements, int on);
/* Encode the digit planes of a device operand AHEAD of the fricc_contract / fricc_contract_pair call that will read
 * it, on the library's current stream (fricc_set_stream) -- e.g. a second stream while T2 is still being uploaded and
 * compressed: the ladder operand Wvvvv does not depend on T2 (fricc/friccsd.py:207 builds it from t1 and the
 * integrals).  The next call that takes the tensor-core form on the same operand (address, shape, strides, term axes,
 * orientation given by out_strides) uses these planes instead of building its own -- once; the caller orders the two
 * streams (the contraction must wait for the preparation) and must not write W in between.  A call that takes the
 * fp64 form drops the preparation.  Results are bit-identical with and without it.  Stat "prepared_plane_hits".
 * An extension of the plugin surface (py_rccsd.prepare_operand).                                                    */
int fricc_prepare_operand(const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                          const int64_t out_strides[4], const char* term);
/* "d_parallel": -1 (default) = the parallel D-selection for candidate lists >= 32768 and the one-warp walk below,
 * 0 / 1 = always serial / always parallel (bit-identical).
 * "pivotal_parallel": -1 (default) = the parallel exact pivotal sampler for N >= 65536 and the one-thread chain kernel
 * below, 0 = always the chain kernel, 1 = always the parallel sampler (bit-identical results).
 * fricc_get_stat: counters of the last call -- "pivotal_restarts" (how many times the parallel sampler had to restart
 * behind a run whose predicted start did not match the chain; 0 in the common case); "tc_contract_calls" /
 * "spmm_contract_calls": how many fricc_contract calls since load took the tensor-core form / the fp64 SpMM.      */
int fricc_get_stat(const char* key, int64_t* value);
/* number of kernels this library has launched since load (bench.py's "gpu_launches" evidence) */
int64_t fricc_kernel_launches(void);

/* ---- uniform draws ------------------------------------------------------------------------------------------
 * The reference draws from libc rand() (fri_utils.cpp:47,74): one draw per systematic compression, one per visited
 * element for pivotal.  The library keeps an equivalent process-wide stream: glibc's TYPE_3 additive-feedback
 * generator, unseeded == srand(1), advanced by exactly the number of draws the reference would have consumed.   */
int fricc_srand(uint32_t seed);                                  /* == libc srand(seed)                           */
int fricc_rand_fill(int32_t* out_host, int64_t n);               /* next n values of rand(); advances the stream  */
int64_t fricc_rand_consumed(void);                               /* draws consumed since the last fricc_srand     */
int fricc_rand_skip(int64_t n);                                  /* advance the stream by n draws (O(log n) jump)  */
/* the next n values of rand() generated ON THE DEVICE into out_dev (per-thread jump-ahead over the linear recurrence);
 * does not advance the stream -- pair it with fricc_rand_skip(consumed)                                          */
/* snapshot / restore of the rand() stream (ring + position), fricc_rand_state_size() bytes: lets ranks that compute
 * different compressions of one iteration keep the single global stream the reference draws from (dist.compress_pair) */
int fricc_rand_state_size(void);
int fricc_rand_get_state(void* buf);
int fricc_rand_set_state(const void* buf);
int fricc_rand_fill_device(int32_t* out_dev, int64_t n);

/* ---- SparseTensor4d (reference: include/sparse_tensor.hpp:13-100, src/py_rccsd.cpp:32-47) -------------------- */
/* ctor #1 (sparse_tensor.hpp:27-31): nnz zero entries */
int fricc_st_create_empty(const int64_t dims[4], int64_t nnz, fricc_st_t* out);
/* ctor #2 (sparse_tensor.hpp:33-77): compress a dense C-contiguous tensor of prod(dims) doubles to m entries.
 * compression: "fri" | "largest"; sampling_method: "pivotal" | "systematic"; verbose: print the 5-row stage table
 * (fri_utils.cpp:509-519).  Draws come from the library stream (above).                                          */
int fricc_st_create(const double* tensor, int mem, const int64_t dims[4], int64_t m, const char* compression,
                    const char* sampling_method, int verbose, fricc_st_t* out);
/* same, with caller-supplied integer draws in [0, 2^31-1] (host memory; draws[t] is what the t-th rand() call would
 * return); n_consumed (may be NULL) receives how many the reference would have consumed.                         */
int fricc_st_create_with_draws(const double* tensor, int mem, const int64_t dims[4], int64_t m,
                               const char* compression, const char* sampling_method, int verbose,
                               const int32_t* draws, int64_t n_draws, int64_t* n_consumed, fricc_st_t* out);
/* build directly from m (flat index, value) pairs (host or device) -- used by the multi-GPU path after the gather  */
int fricc_st_create_from_coo(const int64_t* flat_idx, const double* vals, int mem, const int64_t dims[4], int64_t m,
                             fricc_st_t* out);
/* Provenance hint for the tensor-core form: the first n_exact stored entries are the exactly preserved set D of an FRI
 * compression (include/sparse_tensor.hpp:58-67 stores [D..., S...]); every entry behind them is a sampled one and
 * carries the common magnitude r/n'.  -1 = unknown (from_coo / set_element); nnz for "largest".  A tensor rebuilt from
 * exported entries on another rank carries the hint over with set_n_exact.                                           */
int fricc_st_get_n_exact(fricc_st_t T, int64_t* n_exact);
int fricc_st_set_n_exact(fricc_st_t T, int64_t n_exact);
int fricc_st_destroy(fricc_st_t t);
int fricc_st_size(fricc_st_t t, int64_t* nnz);                                  /* size()       :87              */
int fricc_st_dims(fricc_st_t t, int64_t dims[4]);                               /* dimensions() :86              */
int fricc_st_get_element(fricc_st_t t, int64_t mi, int64_t idx[4], double* value); /* get_element  :138-143        */
int fricc_st_set_element(fricc_st_t t, int64_t mi, int64_t flat_idx, double value); /* set_element  :127-136        */
/* all entries in storage order: flat indices and values (host or device destination)                              */
int fricc_st_export(fricc_st_t t, int64_t* flat_idx, double* vals, int mem);

/* ---- contractions (reference: contract_SparseTensor4d_wrapper, src/sparse_tensor.cpp:349-382) -----------------
 * out[...] += scale * sum_K T[R,K] * W[...]   for term in
 *   "0101" "2323" "1302" "1202" "1303" "1203" "1323" "0112" "0313" "0113" "0312"   (table in SURVEY.md 8a-3)
 * plus the experimental 2323 variants "taskloop" / "atomic" (sparse_tensor.cpp:445-459), which compute the same sum.
 * W and out are 4-D binary64 arrays with arbitrary element strides (numpy semantics); out is accumulated in place.
 * This one call picks the kernel (option "tc_mode"): the bucketed fp64 SpMM, or -- for an FRI-compressed T on a product
 * large enough -- the tensor-core form, with the int8 digit planes expanded straight from the 4-D operand where it
 * lies (no [K][L] fp64 copy), so a driver that calls contract_DTSpT gets the same kernel the sharded engine runs.
 * With every array on the device the call is asynchronous on the library stream.                                    */
int fricc_contract(const double* W, int w_mem, const int64_t w_shape[4], const int64_t w_strides[4], fricc_st_t T,
                   double* out, int out_mem, const int64_t out_shape[4], const int64_t out_strides[4],
                   const char* term);
/* Pre-permuted operand (the multi-GPU column shard): Wp is already the [K][ldw] row-major panel the kernel streams and
 * holds the L-columns [l_begin, l_begin+l_count) of the term's free pair in its first l_count columns.
 * out_panel is [R][ld_out] row-major and receives  out_panel[R][l - l_begin] += scale * sum_K T[R,K] * Wp[K][l - l_begin]. */
int fricc_contract_panel(const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count, fricc_st_t T,
                         double* out_panel_dev, int64_t ld_out, const char* term);
/* Same operand, but accumulated straight into the 4-D strided device output (the single-GPU form of the sharded step:
 * no intermediate panel, no gather):  out[..R.., ..l..] += scale * sum_K T[R,K] * Wp[K][l - l_begin]  for the
 * L-columns l in [l_begin, l_begin + l_count).  (reference: the same terms of src/sparse_tensor.cpp:349-382)          */
int fricc_contract_panel_into(const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count, fricc_st_t T,
                              double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4],
                              const char* term);
/* out[...] += panels[g][R][c] for the n_panels gathered column shards of `term` (panel g holds the L-columns
 * [g*ld, min((g+1)*ld, nL)); panels is [n_panels][nR][ld] row-major on the device, out a 4-D strided device array).   */
int fricc_accumulate_panels(const double* panels_dev, int64_t n_panels, int64_t ld, fricc_st_t T, const char* term,
                            double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4]);
/* ---- exchange over NVLink peer memory (fricc_b200/csrc/fricc_peer.cu) -----------------------------------------
 * The sharded step's one exchange per output, without NCCL on the data path: each rank accumulates its L-columns into
 * its own copy of the 4-D output (fricc_contract_panel_into[_tc]) and then stores exactly those elements into the same
 * addresses of every peer's copy.  The copies live in one symmetric allocation per rank: fricc_symm_alloc returns the
 * device pointer and a 64-byte CUDA IPC handle; peers map it with fricc_symm_open.  (The reference has no counterpart:
 * it is a single-process OpenMP code; the outputs are the `ouput` arrays of contract_DTSpT, src/py_rccsd.cpp:47-50.)   */
int fricc_symm_alloc(int64_t bytes, void** ptr, void* handle64);
int fricc_symm_open(const void* handle64, void** ptr);
int fricc_symm_close(void* ptr);
int fricc_symm_free(void* ptr);
/* peer_out_dev[q] = address of the SAME 4-D output in rank q's copy (q = self: this rank's own, not written);
 * pushes out_local[.., columns l_begin .. l_begin+l_count of the free pair of `term`, ..] to every peer.            */
int fricc_push_columns(const double* out_local_dev, void* const* peer_out_dev, int n_peers, int self, const char* term,
                       const int64_t t_dims[4], const int64_t out_shape[4], const int64_t out_strides[4],
                       int64_t l_begin, int64_t l_count);
/* the same exchange for a 4-D strided VIEW (shape, element strides) of this rank's copy of an output -- the L-slice the
 * sharded engine handed to fricc_contract: every element of the view is stored at the same offset of
 * view_peer_dev[q] (q != self), the address of the same view in peer q's copy.  A store kernel by default; with
 * "push_copy_engine" = 1 long-row views go through the copy engines (stat "push_copies" counts the copies issued).  */
int fricc_push_view(const double* view_local_dev, void* const* view_peer_dev, int n_peers, int self,
                    const int64_t shape[4], const int64_t strides[4]);
/* Stream-ordered barrier of the ranks of one node without an NCCL launch: flags_peer_dev[q] = address, in rank q's
 * symmetric allocation, of 16 zero-initialised 64-bit counters (the same offset on every rank).  Barrier number `epoch`
 * (1, 2, 3, ... -- the same sequence on every rank): one small kernel stores the epoch into every peer's counter for
 * this rank (release: this rank's earlier peer stores are ordered before it) and waits for all peers' epochs in its own
 * counters.  Work queued on the library stream behind the call starts when every rank has reached it.  One process per
 * GPU only: two ranks spinning on ONE device can starve each other.  A rank that never arrives: after option
 * "peer_barrier_timeout_s" seconds (default 300; 0 = wait for ever) the wait gives up WITHOUT trapping (the context
 * survives), records the epoch, and the next fricc_stream_synchronize returns FRICC_ECUDA naming it
 * (fricc_get_stat("peer_barrier_timeouts") reads and clears the same word).                                            */
int fricc_peer_barrier(void* const* flags_peer_dev, int n_peers, int self, uint64_t epoch);
/* synthetic operand for benchmarks: dst[i] = U(0,1) as a function of (seed, first_index + i) only (counter-based), so
 * any sharding of a tensor filled this way holds the same numbers                                                    */
int fricc_fill_uniform(double* dst_dev, int64_t n, uint64_t first_index, uint64_t seed);
/* out_local[.., l, ..] += panel[R][l - l_begin] for this rank's columns (a [nR][ld] panel written by
 * fricc_contract_panel[_tc]); the sums are also stored into every peer's copy when peer_out_dev != NULL.  Transposes
 * through shared memory when out's unit-stride axis is the R-major or L-major index (terms 1302/1202/1303/1203).       */
int fricc_scatter_panel(const double* panel_dev, int64_t ld, int64_t l_begin, int64_t l_count, const char* term,
                        const int64_t t_dims[4], double* out_local_dev, void* const* peer_out_dev, int n_peers, int self,
                        const int64_t out_shape[4], const int64_t out_strides[4]);
/* ---- tensor-core form of the contractions (fricc_b200/csrc/fricc_dense.cu) ------------------------------------
 * Replaces the same reference kernels as fricc_contract_panel (src/sparse_tensor.cpp:12-346) when T comes out of an
 * FRI compression: every sampled entry then has the same magnitude c (fri_utils.cpp:375,502-504), so
 *   out += scale * ( c * SIGN[R,K] x W[K,l]  +  REST[R,K] x W[K,l] ).
 * W is expanded error-free into `n_planes` (1..7) 8-bit digit planes (radix 256, one power-of-two scale per column l);
 * SIGN x planes runs on the int8 tensor cores (tcgen05.mma.kind::i8, exact int32 accumulation), the planes are
 * recombined in fp64, REST (the exactly preserved set D, duplicates, any entry whose magnitude is not c) goes through
 * the fp64 SpMM.  The only approximation is the rounded tail of the expansion: <= 2^(-8*n_planes) of the column's
 * largest |W| per term (n_planes = 6: 2^-48).                                                                        */
typedef struct fricc_wslices* fricc_wslices_t; /* digit planes of one W panel (device) */
/* planes of the panel columns [0, l_count) of Wp_dev[nK][ldw] */
int fricc_wslices_create(const double* Wp_dev, int64_t ldw, int64_t nK, int64_t l_count, int n_planes,
                         fricc_wslices_t* out);
/* planes of the columns [l_begin, l_begin + l_count) of the [K][L] operand of `term`, read where it lies inside the 4-D
 * device array W (element strides): no fp64 panel is formed.  Operands whose contracted run is contiguous -- the
 * ladder's native W[a,b,c,d] -- go through a cluster kernel that reads every element from HBM once.                  */
int fricc_wslices_create_view(const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4], const char* term,
                              int64_t l_begin, int64_t l_count, int n_planes, fricc_wslices_t* out);
/* re-expand in place (the operand changed, same shape and column count): what a driver that rebuilds W every iteration
 * calls (fricc/friccsd.py:207,325-331)                                                                               */
int fricc_wslices_update_view(fricc_wslices_t h, const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                              const char* term, int64_t l_begin);
/* An update of planes that exist already re-uses their column exponents as a prediction (option "tc_predict", default
 * 1): one streaming pass cuts the digits and takes the new column maxima, and only the columns whose exponent moved are
 * cut again -- bit-identical to a fresh build.  How many columns that was in the last update (-1: the last build was
 * not a predicted one).  Reads 4 bytes back from the device.                                                          */
int fricc_wslices_redone(fricc_wslices_t h, int64_t* n_columns);
/* measured int8 tensor-core peak of the current device: back-to-back tcgen05.mma.kind::i8 (M=128, N=256, K=32) on
 * operands resident in shared memory, one CTA per SM, for about `seconds` -> int8 TOP/s (2 x MAC).  bench.py's
 * denominator for the tensor-core kernel's roofline (short run = burst clocks, seconds = under the power cap).        */
int fricc_int8_peak(double seconds, double* tops);
int fricc_wslices_update(fricc_wslices_t h, const double* Wp_dev, int64_t ldw); /* re-expand (W changed), same shape */
/* out[.., l, ..] += scale * sum_K T[R,K] * W[K,l] for the columns l in [l_begin, l_begin + planes' l_count) with planes
 * the caller keeps (an operand that does not change between iterations: built once with fricc_wslices_create_view);
 * W_dev is the same 4-D operand, read where it lies for the exact part of T.  (fricc_contract does both steps in one
 * call; this is the two-step form for drivers that hold on to the planes.)                                           */
int fricc_contract_planes(fricc_wslices_t h, const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                          int64_t l_begin, fricc_st_t T, double* out_dev, const int64_t out_shape[4],
                          const int64_t out_strides[4], const char* term);
int fricc_wslices_destroy(fricc_wslices_t h);
int fricc_wslices_bytes(fricc_wslices_t h, int64_t* bytes);
/* Wrec_dev[nK][l_count] = the fp64 value the planes represent (for tests of the expansion) */
int fricc_wslices_reconstruct(fricc_wslices_t h, double* Wrec_dev);
/* fraction of T's entries that the layout of `term` holds in the sign matrix (0 for "largest" compressions) */
int fricc_st_sign_fraction(fricc_st_t T, const char* term, double* fraction);
/* same contract as fricc_contract_panel / fricc_contract_panel_into; l_count is the planes' l_count.
 * overwrite != 0: out_panel[R][0 .. l_count) = result instead of += (a fresh panel needs neither clearing nor reading) */
int fricc_contract_panel_tc(fricc_wslices_t h, const double* Wp_dev, int64_t ldw, int64_t l_begin, fricc_st_t T,
                            double* out_panel_dev, int64_t ld_out, const char* term, int overwrite);
int fricc_contract_panel_into_tc(fricc_wslices_t h, const double* Wp_dev, int64_t ldw, int64_t l_begin, fricc_st_t T,
                                 double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4],
                                 const char* term);
/* permute a 4-D W into the [K][L] panel layout of `term` on the device (columns [l_begin, l_begin+l_count))        */
int fricc_permute_w(const double* W, int w_mem, const int64_t w_shape[4], const int64_t w_strides[4],
                    const char* term, int64_t l_begin, int64_t l_count, double* Wp_dev);

/* ---- stage-wise FRI utilities (reference: include/fri_utils.hpp:8-32) ------------------------------------------ */
/* one_norm (fri_utils.cpp:164-171): sum |x_i| accumulated in index order (bit-identical to the 1-thread reference) */
int fricc_one_norm(const double* x, int mem, int64_t n, double* norm);
/* The same sum for callers that must not allocate device memory while they run.  NO COUNTERPART IN THE REFERENCE (a host
 * code; its one_norm has no working memory at all).  Meant for a driver whose own arena already fills the GPU, for a
 * host thread that cannot afford a trip into the kernel driver (an allocation that maps memory can stall for seconds
 * behind other processes' queries), and for processes held to a fixed device-memory budget.  The chunk sums, chunk maps
 * and carries of the exact scan live in a fixed 1.3 MB array inside the library image (present on every GPU that loads
 * the library, until exit) instead of the block cache; the call issues kernels, two small memsets and -- for a host
 * result -- one 8-byte copy, and reaches no cudaMalloc/cudaFree variant and no allocating routine.
 *   x_dev    device pointer (a host array would need staging, i.e. an allocation), n <= FRICC_ONE_NORM_STATIC_MAX_N;
 *            larger n -> FRICC_EINVAL with a message that names the limit
 *   norm     where the sum goes: norm_mem = FRICC_DEVICE -> written by the last kernel, the call stays asynchronous on
 *            the library stream; FRICC_HOST -> copied back, the stream is synchronised
 * Bit-identical to fricc_one_norm.  All calls on a GPU share the one array and it keeps the previous call's contents:
 * what has to start at zero is cleared by every call, and calls are safe one after another on one stream, but two calls
 * that can overlap on the GPU (two streams without an ordering, two host threads) race on it.  The first call on each
 * device looks up the array's address there (cudaGetSymbolAddress under the current device) and keeps it: make that
 * first call during set-up, before any stream capture is started.                                                     */
#define FRICC_ONE_NORM_STATIC_MAX_N (1 << 26)
int fricc_one_norm_static(const double* x_dev, int64_t n, double* norm, int norm_mem);
/* argsort (fri_utils.cpp:11-36): indices of the m largest |x| in descending |x| (ties: ascending index)            */
int fricc_argsort(const double* x, int mem, int64_t n, int64_t m, int64_t* idx_out_host);
/* get_d_largest (fri_utils.cpp:400-437) */
int fricc_get_d_largest(const double* x, int mem, int64_t n, int64_t n_sample, int64_t* D_out_host, int64_t* n_D,
                        double* remaining_norm);
/* make_probability_vector (fri_utils.cpp:367-398) */
int fricc_make_probability_vector(const double* x, int mem, int64_t n, int64_t n_sample, const int64_t* D_host,
                                  int64_t n_D, double remaining_norm, double* p_out, int p_mem);
/* sample_pivotal (fri_utils.cpp:63-127) / sample_systematic (:129-159) on an explicit probability vector          */
int fricc_sample_pivotal(int64_t n_sample, const double* probs, int mem, int64_t n, const int32_t* draws,
                         int64_t n_draws, int64_t* n_consumed, int64_t* S_out_host);
int fricc_sample_systematic(int64_t n_sample, const double* probs, int mem, int64_t n, const int32_t* draws,
                            int64_t n_draws, int64_t* n_consumed, int64_t* S_out_host);
/* parallel_sample (fri_utils.cpp:176-362, "Alg. A.2"): probs cut into n_chunks contiguous chunks (the reference uses
 * omp_get_max_threads()), integer budgets by a pivotal draw over the fractional parts, leading probabilities of each
 * chunk rescaled to its budget, chunks sampled independently with sample_pivotal.  The reference's threads call rand()
 * concurrently (a race); here the draws are consumed in chunk order -- its behaviour when the threads run one after
 * another, exactly its behaviour with one chunk.                                                                    */
int fricc_parallel_sample(int64_t n_sample, const double* probs, int mem, int64_t n, int n_chunks, const int32_t* draws,
                          int64_t n_draws, int64_t* n_consumed, int64_t* S_out_host);
/* fri_compression (fri_utils.cpp:439-522): returns m flat indices and m values, [D..., S...] order               */
int fricc_fri_compression(const double* x, int mem, int64_t n, int64_t n_sample, const char* sampling_method,
                          int verbose, const int32_t* draws, int64_t n_draws, int64_t* n_consumed,
                          int64_t* idx_out_host, double* vals_out_host);

/* ---- range-sharded compression (multi-GPU: x sharded by contiguous flat-index range; fricc_b200/csrc/fricc_shard.cu) --
 * Rank g holds x[j0, j0+n) of an n_total-element tensor; rank order = index order.  These are the per-shard stages of
 * fri_compression (fri_utils.cpp:439-522) between which the host exchanges a few bytes (fricc_b200/dist.py:
 * compress_ranges); the assembled result is bit-identical to the single-process compression.  All *_dev scalars live on
 * the device, so a chain of carries from rank to rank is ordered by the stream alone.                               */
/* 4096-bin histogram (count) of the 12-bit digit at `shift` (51,39,27,15,3; shift 0 = the last 3 bits) of
 * key(x)=bits(|x|), over elements whose higher bits equal `prefix` (prefix_bits > 0 enables the filter); ADDS into the
 * device array, which the caller all-reduces (argsort, fri_utils.cpp:11-36, without the sort).                      */
int fricc_key_histogram(const double* x_dev, int64_t n, int shift, uint64_t prefix, int prefix_bits,
                        unsigned long long* hist_dev /*[4096]*/);
typedef struct fricc_shard* fricc_shard_t;
int fricc_shard_create(const double* x_dev, int64_t n, int64_t j0, int64_t n_total, fricc_shard_t* out);
int fricc_shard_destroy(fricc_shard_t s);
/* what the exact running sum adds up: |x_j| (one_norm, fri_utils.cpp:164-171) ...                                    */
int fricc_shard_use_abs(fricc_shard_t s);
/* ... or p_j = |x_j| / remaining_norm * n_sample with p = 0 on D (make_probability_vector, fri_utils.cpp:367-398);
 * D = { key > key_cut } U { key == key_cut and flat index <= idx_cut }; no D: key_cut = ~0                           */
int fricc_shard_use_probabilities(fricc_shard_t s, double remaining_norm, int64_t n_sample, uint64_t key_cut,
                                  uint64_t idx_cut);
/* exact LEFT-TO-RIGHT sum across shards in three stages (std::partial_sum / one_norm order):
 * (1) approximate sum of this shard -> *approx_total_dev; (2) chunk maps, given the approximate sum of all lower shards;
 * (3) exact: *carry_out_dev = exact sum through this shard, given the exact sum *carry_in_dev through the lower ones. */
int fricc_shard_scan_local(fricc_shard_t s, double* approx_total_dev);
int fricc_shard_scan_maps(fricc_shard_t s, const double* approx_carry_dev);
int fricc_shard_scan_resolve(fricc_shard_t s, const double* carry_in_dev, double* carry_out_dev);
/* sample_systematic (fri_utils.cpp:129-159) after stage (3) over the probabilities: the samples owned by this shard's
 * elements, as GLOBAL flat indices, into S_dev[n_sample] (zero-initialised by the caller; summed over ranks).
 * draw = the rand() value of lin_space_add_const (fri_utils.cpp:47).                                                */
int fricc_shard_emit_systematic(fricc_shard_t s, int64_t n_sample, int32_t draw, int64_t* S_dev);
/* The same for n_draws consecutive compressions of the same tensor from ONE pass (fricc/friccsd.py:100-136 compresses
 * the same t2 twice; the running sums of p do not depend on the draw): draw d fills S_dev + d * stride.  Identical to
 * n_draws calls of fricc_shard_emit_systematic.                                                                     */
int fricc_shard_emit_systematic_multi(fricc_shard_t s, int64_t n_sample, const int32_t* draws, int n_draws, int64_t* S_dev,
                                      int64_t stride);
/* Alg. A.2 (parallel_sample, fri_utils.cpp:176-362) with ONE CHUNK PER RANK -- the order-independent distributed sampler
 * of SURVEY.md 8 f-4.  The shard is the chunk; the caller has done the budget arithmetic of :191-275 (exact |q|_1 over
 * all ranks and |q_j|_1 of this chunk with the scan stages above, a_j = n_sample*|q_j|/|q|, t_j = frac(a_j), the integer
 * budget g_j after the pivotal draw over the t's).  This call is :296-356 for the chunk: copy of the probabilities,
 * leading ones rescaled to the mass `budget`, sample_pivotal, indices offset by j0 -> S_dev[0 .. budget).  draws: this
 * chunk's own uniform draws (host int32, rand() values), or NULL = the library's rand() stream at its current position
 * (the caller positions each rank's copy with fricc_rand_skip).  n_consumed: draws the chunk's walk took.           */
int fricc_shard_sample_chunk(fricc_shard_t s, int64_t budget, double t, double a, double chunk_norm, const int32_t* draws,
                             int64_t n_draws, int64_t* n_consumed, int64_t* S_dev);
/* raises the reference's error if a kernel of this shard flagged one (some p > 1: fri_utils.cpp:383-395)            */
int fricc_shard_check(fricc_shard_t s);
/* (key, global flat index) of every element with key >= key_min, appended in arbitrary order; *count_dev = how many
 * (may exceed cap: then only cap were stored)                                                                       */
int fricc_shard_candidates(fricc_shard_t s, uint64_t key_min, uint64_t* keys_out_dev, uint64_t* idx_out_dev, int64_t cap,
                           unsigned long long* count_dev);
/* order gathered candidates by (key descending, index ascending), in place                                          */
int fricc_sort_candidates(uint64_t* keys_dev, uint64_t* idx_dev, int64_t count, int64_t n_total);
/* get_d_largest's greedy test (fri_utils.cpp:415-431) on an ordered candidate list                                   */
int fricc_d_greedy(const uint64_t* keys_sorted_dev, int64_t n_candidates, int64_t n_sample, double norm, int64_t* n_D,
                   double* remaining_norm);
/* values (fri_utils.cpp:498-504) of the entries whose flat index this shard owns: x for the first n_D entries, x/p for
 * the samples; 0.0 for entries owned by another shard (sum over ranks assembles the list)                           */
int fricc_shard_values(fricc_shard_t s, const int64_t* flat_dev, int64_t m, int64_t n_D, double* vals_dev);

#ifdef __cplusplus
}
#endif
#endif /* FRICC_B200_H */
