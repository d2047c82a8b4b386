/*
 * ckb.h -- C ABI of the B200 (sm_100a) mixed-precision GEMM library (libckb.so).
 *
 * This is the drop-in boundary below Composable Kernel's device-operator layer.  CK's host classes
 * (include/ck/tensor_operation/gpu/device/...) keep their names and signatures; where the reference
 * launches a HIP `__global__` kernel through launch_and_time_kernel
 * (include/ck/host_utility/kernel_launch.hpp:13-84) our classes call the entry points below instead.
 *
 *   reference interface                                                   -> entry point here
 *   --------------------------------------------------------------------------------------------
 *   DeviceFpAintBGemm_Wmma_CShuffle::{Argument ctor, IsSupportedArgument,
 *     Invoker::Run}   device/impl/device_fpAintB_gemm_wmma.hpp:345-394,424-494,510-591
 *                                                                          ckb_gemm_query / ckb_gemm_run
 *   DeviceGemmMultipleABD_Xdl_CShuffle::{Argument, IsSupportedArgument, Invoker::Run}
 *     device/impl/device_gemm_multiple_abd_xdl_cshuffle.hpp:91-138,145-571,587-603
 *                                                                          ckb_gemm_query / ckb_gemm_run
 *   DeviceGemmMultipleD_Xdl_CShuffle (int8 x int8, quantization epilogues)
 *     device/impl/device_gemm_multiple_d_xdl_cshuffle.hpp                  ckb_gemm_query / ckb_gemm_run
 *   DeviceGemmMultiD_Xdl_CShuffle_V3 (+MultiplyMultiply)
 *     device/impl/device_gemm_multiple_d_xdl_cshuffle_v3.hpp:143-175       ckb_gemm_query / ckb_gemm_run
 *   DeviceGemm_Xdl_CShuffleV3R1 (split-K + reduce, GetWorkSpaceSize)
 *     device/impl/device_gemm_xdl_cshuffle_v3r1.hpp:188-294,686-694        ckb_gemm_query (workspace_bytes)
 *   DeviceGroupedGemm_Xdl_Multi_ABD_Fixed_NK::{SetDeviceKernelArgs, Run}
 *     device/impl/device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:37-145,580-676
 *                                                                          ckb_grouped_gemm_run
 *   UnsignedWeightPreprocessor  example/64_fpAintB_gemm/common.hpp:56-83   ckb_prepack_b_u8
 *   [EXT] packed-int4 weight pre-pack (no reference counterpart)           ckb_prepack_b_int4
 *
 * Conventions: plain pointers and sizes, no C++ or torch types.  All data pointers are DEVICE
 * pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.  Strides are in
 * ELEMENTS (ck::index_t semantics).  Every function returns 0 on success and a negative ckb_status
 * otherwise; ckb_last_error() gives the thread-local message.  There is no CPU fallback: without a
 * CUDA device of compute capability 10.x every run call fails with CKB_ERR_NO_DEVICE.
 */
#ifndef CKB_H
#define CKB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CKB_VERSION 100

typedef enum ckb_status {
    CKB_OK              = 0,
    CKB_ERR_UNSUPPORTED = -1, /* IsSupportedArgument == false */
    CKB_ERR_INVALID     = -2, /* null pointer / bad enum / negative size */
    CKB_ERR_NO_DEVICE   = -3, /* no sm_100 device / driver entry point missing */
    CKB_ERR_CUDA        = -4, /* a CUDA call failed (message in ckb_last_error) */
    CKB_ERR_WORKSPACE   = -5  /* workspace required but not supplied */
} ckb_status;

typedef enum ckb_dtype {
    CKB_F32 = 0, CKB_F16 = 1, CKB_BF16 = 2, CKB_I8 = 3, CKB_U8 = 4, CKB_I32 = 5,
    CKB_I4  = 6 /* two 4-bit codes per byte; only in the pre-packed layout of ckb_prepack_b_int4 */
} ckb_dtype;

typedef enum ckb_layout {
    CKB_ROW = 0,      /* tensor_layout::gemm::RowMajor    */
    CKB_COL = 1,      /* tensor_layout::gemm::ColumnMajor */
    CKB_PREPACKED = 2 /* B only: output of ckb_prepack_b_int4 */
} ckb_layout;

/* B element-wise op: how the stored weight becomes the GEMM operand */
typedef enum ckb_b_op {
    CKB_BOP_PASSTHROUGH = 0, /* PassThrough<AType,int8>                 unary_element_wise_operation.hpp:112-121 */
    CKB_BOP_MULTIPLY    = 1, /* Multiply<bhalf_t,int8_t,bhalf_t>(B0,B1)  binary_element_wise_operation.hpp:184-190 */
    CKB_BOP_FPAINTB     = 2, /* fp16(u8 - 128) * scale[n]                reference_fpAintB_gemm.hpp:110, B = biased uint8 */
    CKB_BOP_GROUP       = 3, /* [EXT] W4A16 group scale, ab_scale numerics: exact (q - zp[k/g,n]) in A's type into the MMA, the group
                                scale s[k/g,n] multiplies the fp32 partial sum of each group (host pattern:
                                gemm_multiply_multiply_xdl_fp8_ab_scale.cpp:258-290; device_gemm_multiple_d_ab_scale.hpp:22-59).
                                Every M: HMMA decode kernels up to 64 rows per launch, row slabs of them above. */
    CKB_BOP_GROUP_RTN   = 4  /* [EXT] W4A16 group scale, dequant-in-B-copy numerics: (q - zp) * s rounded ONCE (nearest-even) into
                                A's type, then a plain 16-bit GEMM (the weight threadwise_tensor_slice_transfer_v3r1_dequant.hpp
                                :509-547 hands the block GEMM).  Every M: fused tcgen05 kernel above 64 rows, HMMA kernel below. */
} ckb_b_op;

/* CDE element-wise op (epilogue).  Same numbering as the oracle's orc_cde_op. */
typedef enum ckb_cde_op {
    CKB_CDE_PASSTHROUGH           = 0,
    CKB_CDE_ADD                   = 1,
    CKB_CDE_FASTGELU              = 2,
    CKB_CDE_ADD_FASTGELU          = 3,
    CKB_CDE_MULTIPLY              = 4,
    CKB_CDE_MULTIPLY_ADD          = 5,
    CKB_CDE_MULTIPLY_FASTGELU     = 6,
    CKB_CDE_MULTIPLY_ADD_FASTGELU = 7,
    CKB_CDE_MULTIPLY_MULTIPLY     = 8,
    CKB_CDE_ACT_MUL_CLAMP         = 9,
    CKB_CDE_ADD_ACT_MUL_CLAMP     = 10,
    CKB_CDE_ACT_MUL2_CLAMP        = 11,
    CKB_CDE_ADD_ACT_MUL2_CLAMP    = 12
} ckb_cde_op;

typedef enum ckb_act { CKB_ACT_PASSTHROUGH = 0, CKB_ACT_RELU = 1 } ckb_act;

enum {
    CKB_FLAG_BF16_RNE = 1 << 0, /* round bf16 outputs to nearest-even instead of CK's truncation
                                  (type_convert.hpp:53-64 vs :508-549); default = truncation */
    CKB_FLAG_FIXED_K_ORDER = 1 << 1, /* [EXT] W4A16 decode (M <= 64): make the fp32 summation order over K independent of N, of the
                                  grid and of the shard a problem is cut into, so that G column shards + all-gather reproduce
                                  the single-GPU E bit for bit.  Without it the planner picks the fastest warp
                                  grid per shape and two different N may differ in the last bit of a few outputs. */
    /* kernel-choice overrides: the library picks per shape by default; the instance factories register the alternatives as
       separate instances so that a profiler can time them (the reference's run-every-instance-keep-the-best loop) */
    CKB_FLAG_SS_1CTA   = 1 << 2,            /* dense / int8 tcgen05 kernel: one CTA per 128x256 tile */
    CKB_FLAG_SS_2CTA   = 1 << 3,            /* ... CTA pair (cta_group::2) per 256x256 tile */
    CKB_FLAG_PREFILL_TWO_PASS = 1 << 4,     /* weight-quantized prefill: dequant pass into workspace + dense GEMM (round-1 form) */
    CKB_FLAG_GEMV_SINGLE_PRODUCER = 1 << 5, /* W4A16 decode: single-producer kernel (wq_gemv) instead of the rotating-producer one */
    CKB_FLAG_W8_STREAMK = 1 << 7,           /* int8-weight decode rows: the stream-K tcgen05 kernel (TMEM A operand) instead of the HMMA one */
    CKB_FLAG_INPUTS_READY = 1 << 6          /* [EXT] the caller asserts that A, D and the weights of THIS GEMM are not written by the
                                               kernel launched just before it on the stream (e.g. the k / v projections after q, `up` after
                                               `gate`: same layer input).  Under programmatic dependent launch (graph replay) the W4A16 decode
                                               kernel then starts computing while its predecessor drains, and waits for the predecessor only
                                               before it exits, so completion order along the stream is preserved.  No effect elsewhere. */
};

#define CKB_MAX_D 2
#define CKB_MAX_PEERS 8

/* E[M,N] = cde( A[M,K] * bop(B[K,N], scale[, zero]), D0, D1 ) */
typedef struct ckb_gemm_problem {
    int32_t M, N, K;

    int32_t a_dtype;  /* CKB_F16 | CKB_BF16 | CKB_I8 */
    int32_t a_layout; /* CKB_ROW (A[m,k] = a[m*lda + k]) */
    const void* a;
    int64_t lda;

    int32_t b_dtype;  /* CKB_I8 | CKB_U8 (biased, +128) | CKB_I4 */
    int32_t b_layout; /* CKB_COL: B[k,n] = b[n*ldb + k];  CKB_ROW: b[k*ldb + n];  CKB_PREPACKED */
    const void* b;
    int64_t ldb;
    int32_t b_op;

    /* per-channel scale (MULTIPLY, FPAINTB): scale[n], element type = A's type (CK passes it with
       stride 0 along K: StrideBs = {StrideB, 0}).  GROUP: scale[(k/group_size)*N + n]. */
    int32_t s_dtype;
    const void* scale;
    /* GROUP only: zero points, uint8 [(k/group_size)*N + n], or NULL -> zero_const for all */
    const void* zero;
    int32_t zero_const;
    int32_t group_size; /* 128 */

    /* D(m,n) = d[i][m*d_sm[i] + n*d_sn[i]]  (Row layout, StrideD s: {s,1}; Col: {1,s}; s = 0 broadcasts) */
    int32_t num_d;
    int32_t d_dtype[CKB_MAX_D];
    const void* d[CKB_MAX_D];
    int64_t d_sm[CKB_MAX_D], d_sn[CKB_MAX_D];

    int32_t cde_op;
    float cde_scale; /* requantScale_ of Activation_Mul_Clamp / Add_Activation_Mul_Clamp */
    int32_t act;

    int32_t e_dtype; /* CKB_F16 | CKB_BF16 | CKB_F32 | CKB_I8 | CKB_I32 */
    void* e;         /* row-major, E[m,n] = e[m*lde + n] */
    int64_t lde;

    int32_t kbatch; /* split-K hint (KBatch / KSplit of DeviceGemmV2R1); 0 or 1 = library decides */
    uint32_t flags;

    /* [EXT] fused all-gather of the output over NVLink peer memory (column-parallel shards across the GPUs of one box,
       SURVEY.md 8e; the reference has no multi-GPU code).  num_peers = G > 0: the epilogue stores every element of this
       rank's E slice at e_peer[g][m*lde + n] for g = 0..G-1 -- e_peer[g] is THIS rank's slot inside rank g's gathered
       buffer, a peer-mapped device pointer (cudaIpc / cuMem / torch symmetric memory), own rank included; `e` is not
       written.  flag_peer[g] points to TWO uint32 in rank g's memory: [0] the arrival counter, [1] a scratch word only rank
       g's own kernels touch (both start at 0).  When the last CTA of the launch has finished (device-scope count on the own
       scratch word) it adds 1 (release, system scope) to the arrival counter of every rank; a rank's gathered buffer is
       complete when its counter has grown by G (ckb_peer_wait).  peer_rank = this rank's index in the arrays.
       Served by the W4A16 decode kernel (CKB_BOP_GROUP, M <= 32); other problems report unsupported. */
    int32_t num_peers, peer_rank;
    void* e_peer[CKB_MAX_PEERS];
    void* flag_peer[CKB_MAX_PEERS];
} ckb_gemm_problem;

typedef struct ckb_gemm_plan {
    int32_t supported;       /* IsSupportedArgument */
    int32_t kernel_id;       /* which sm_100a kernel serves the problem */
    char kernel_name[96];    /* GetTypeString() tail */
    uint64_t workspace_bytes; /* GetWorkSpaceSize; workspace must be zero-initialised ONCE by the caller
                                 (cudaMemset) -- kernels restore it to zero before they exit */
    int32_t grid, block, smem_bytes;
    int32_t split_units;
    uint64_t algo_bytes;     /* algorithmic HBM bytes of one launch: A + B + scales[/zp] + D + E */
    uint64_t flops;          /* 2*M*N*K */
    /* TMA descriptors etc. encoded at query time for exactly the pointers in the problem */
    unsigned char opaque[1280] __attribute__((aligned(64)));
} ckb_gemm_plan;

/* Validate the problem, choose the kernel, size the workspace, encode the TMA descriptors.
   Returns CKB_OK with plan->supported = 0/1 (unsupported is not an error here). */
int ckb_gemm_query(const ckb_gemm_problem* problem, ckb_gemm_plan* plan);

/* Launch on `stream`.  `plan` must come from ckb_gemm_query on the same problem (same pointers). */
int ckb_gemm_run(const ckb_gemm_problem* problem, const ckb_gemm_plan* plan, void* workspace, void* stream);

/* Allocation-free variant of ckb_gemm_run for the one kernel family that needs scratch memory at decode sizes (the stream-K
   tcgen05 decode kernel, plan->kernel_name "wq_decode_*": int8-weight formats behind CKB_FLAG_W8_STREAMK, M <= 64 rows).
   NO counterpart in the reference: added for callers that cannot call
   cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The scratch (fp32 partial
   tiles + per-tile arrival counters, plan->workspace_bytes) lives in a file-scope `static __device__` array of
   CKB_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds until the process exits.  The call
   allocates nothing (no cudaMalloc / cudaFree / allocating library routine): it clears the part of the array it uses
   (cudaMemsetAsync on `stream`) and launches the same kernel as ckb_gemm_run, so results equal ckb_gemm_run's.
   Limits: plan->workspace_bytes <= CKB_STATIC_SCRATCH_BYTES (8 MiB: N <= 28672 at M <= 64), else CKB_ERR_WORKSPACE; plans of
   other kernels (which need no scratch: use ckb_gemm_run with workspace = NULL) and M > 64 are CKB_ERR_UNSUPPORTED.
   EVERY call on a device shares the one array: calls that may overlap on the GPU (different streams, concurrent graph branches)
   race on it -- keep them on one stream or order them yourself. */
#define CKB_STATIC_SCRATCH_BYTES (8u << 20)
uint64_t ckb_static_scratch_bytes(void);
int ckb_gemm_run_static_scratch(const ckb_gemm_problem* problem, const ckb_gemm_plan* plan, void* stream);

/* Host-buffer convenience used for end-to-end measurement: A, D and E are HOST pointers (pinned or
   pageable); weights/scales stay device-resident.  Copies A (and D) host->device, runs, copies E
   device->host, synchronises the stream.  `staging` = device scratch of ckb_gemm_host_staging_bytes(). */
uint64_t ckb_gemm_host_staging_bytes(const ckb_gemm_problem* problem);
int ckb_gemm_run_host(const ckb_gemm_problem* problem, void* staging, void* workspace, void* stream);
/* Same, without the final synchronisation: copies and kernel are only ENQUEUED on `stream` (A, D, E must be pinned
   host memory that stays valid until the caller synchronises).  Calls on one stream may share one staging buffer:
   stream order keeps GEMM i's E read-back ahead of GEMM i+1's A upload. */
int ckb_gemm_run_host_async(const ckb_gemm_problem* problem, void* staging, void* workspace, void* stream);

/* ---- weight pre-packing (one-time; excluded from timing exactly as the reference excludes
        UnsignedWeightPreprocessor, example/64_fpAintB_gemm/run_gemm_example.inc:52-58) -------- */

/* int8 -> biased uint8 (u8 = int8 + 128), n elements, device pointers. */
int ckb_prepack_b_u8(const void* src_i8, void* dst_u8, int64_t n, void* stream);

/* Row-major int8 B[k,n] = src[k*ld_src + n] (CK's mk_kn_mn) -> K-major dst[n*K + k], the layout the kernels
   stream.  One-time weight transform, device pointers. */
int ckb_prepack_b_rowmajor_i8(const void* src_kn, int64_t ld_src, int32_t K, int32_t N, void* dst_nk, void* stream);

/* Column-major A (CK's km_*: A[m,k] = src[k*ld_src + m], 16-bit or 8-bit elements) -> row-major dst[m*K + k].  Activations change
   every call, so the C++ operators run this on the launch stream in front of every GEMM, into caller-supplied workspace. */
int ckb_transpose_a_colmajor(const void* src_km, int64_t ld_src, int32_t M, int32_t K, int32_t elem_bytes, void* dst_mk, void* stream);

/* [EXT] 4-bit codes, one per byte, logical B[k,n] = src[k*src_sk + n*src_sn] (device) ->
   tile-major packed layout consumed by the W4A16 kernels.  dst needs ckb_prepack_b_int4_bytes(). */
uint64_t ckb_prepack_b_int4_bytes(int32_t N, int32_t K);
int ckb_prepack_b_int4(const void* src_u8, int64_t src_sk, int64_t src_sn, int32_t N, int32_t K,
                       void* dst, void* stream);

/* ---- grouped GEMM with fixed N,K (device-side argument array, client_example/31) ------------- */
typedef struct ckb_group_arg { /* mirrors GroupedGemmMultiABDKernelArgument<1,2,1>
                                  device_grouped_gemm_multi_abd.hpp:15-24 (device-resident array) */
    const void* a;     /* As[0] */
    const void* b0;    /* Bs[0] int8 */
    const void* b1;    /* Bs[1] scale */
    const void* d0;    /* Ds[0] */
    void* e;
    int32_t M, N, K;
    int32_t stride_a, stride_b0, stride_b1, stride_d0, stride_e;
} ckb_group_arg;

/* Size of one element of the device argument array, GroupedGemmMultiABDKernelArgument<1, num_b, num_d>
   (GetDeviceKernelArgSize / group_count).  ckb_group_arg above is the num_b = 2, num_d = 1 case; with fewer
   tensors the pointer and stride slots of the missing ones are simply absent, exactly as in the reference struct. */
uint64_t ckb_grouped_gemm_arg_bytes(int32_t num_b, int32_t num_d);

/* IsSupportedArgument for the grouped op (device_grouped_gemm_multi_abd_xdl_fixed_nk.hpp:580-610).  `proto`
   carries what the host knows: dtypes, layouts, ops, N, K, the strides (lda, ldb, lde), num_d / d_dtype[0];
   b_op = MULTIPLY means Bs = (int8 B0, bf16 B1 per channel), PASSTHROUGH means Bs = (B0).  Pointers and M are
   ignored.  Returns CKB_OK or CKB_ERR_UNSUPPORTED. */
int ckb_grouped_gemm_query(const ckb_gemm_problem* proto, int32_t group_count);

/* Invoker::Run (:613-676).  `dev_args` = DEVICE array of group_count argument structs (SetDeviceKernelArgs);
   M of every group is read on the device; sum_m (gemm_descs[i].M_ == sum of all M, :460-463) sizes the grid. */
int ckb_grouped_gemm_run(const ckb_gemm_problem* proto, const void* dev_args, int32_t group_count,
                         int32_t sum_m, void* stream);

/* [EXT] consumer side of the fused all-gather: enqueue a one-thread kernel on `stream` that advances the epoch counter
   *epoch_state (device uint32, starts at 0) and returns once *flag (this rank's arrival counter, device uint32, starts
   at 0) has reached epoch * arrivals_per_epoch.  Everything enqueued on `stream` after it sees the gathered buffer.
   Graph-capturable (all state lives in device memory). */
int ckb_peer_wait(void* flag, void* epoch_state, uint32_t arrivals_per_epoch, void* stream);

/* ---- scheduler arithmetic, host only (no device needed): the functions the kernels themselves use, exported so that the
        tile maps are unit-tested on the CPU in the style of the reference's test/block_to_ctile_map ------------------------- */
/* output-tile raster of the persistent tcgen05 kernels: linear tile index t -> (m block, n block), bands of `group` M blocks */
int ckb_tile_coord(int32_t t, int32_t num_m_blk, int32_t num_n_blk, int32_t group, int32_t* mb, int32_t* nb);
/* stream-K unit range [begin, end) of CTA `cta` out of `grid` */
int ckb_streamk_range(int64_t total_units, int32_t cta, int32_t grid, int64_t* begin, int64_t* end);
/* W4A16 decode kernel schedule for (M, N, K) on `sms` SMs; mode bit 0 = CKB_FLAG_FIXED_K_ORDER, bit 1 = GROUP_RTN.  out[24] =
   {nt, rb, n_groups, kb_total, grid, passes, n_slots, slot_q, slot_r, g_hi, wn_log2, gw, kbu, upw, n_chunks, ring, stage_bytes,
    off_w, off_s, off_z, red_off, bar_off, smem_bytes, warps}.  Returns CKB_ERR_UNSUPPORTED when no schedule fits. */
int ckb_gemv_schedule(int32_t M, int32_t N, int32_t K, int32_t zero_points, int32_t sms, int32_t mode, int32_t* out);

/* ---- misc --------------------------------------------------------------------------------- */
const char* ckb_last_error(void);
int ckb_version(void);
int ckb_device_sm_count(void); /* <0 when no usable device */

#ifdef __cplusplus
}
#endif
#endif
