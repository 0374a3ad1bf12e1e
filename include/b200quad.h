/* b200quad — Blackwell-native numerical integration: public C ABI.
 *
 * Everything the Python layer, the torch ops and the CLIs do goes through the
 * entry points declared here.  Reference parity notes cite the upstream
 * Excalibur1224/Cuda-v-MPI sources as file:line.
 */
#ifndef B200QUAD_H_
#define B200QUAD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums ------------------------------------------------------------- */

/* Integrand library.  SIN is the hard-coded integrand of riemann.cpp:37 and
 * cintegrate.cu:68; ACC/VEL/DIS are the analytic train model of
 * riemann.cpp:103-116; PROFILE is the 1801-entry velocity table of ex4vel.h
 * lerped as faccel() does (4main.c:262-268, cintegrate.cu:36-44). */
enum bq_integrand {
    BQ_F_SIN = 0,
    BQ_F_ACC = 1,
    BQ_F_VEL = 2,
    BQ_F_DIS = 3,
    BQ_F_PROFILE = 4,
    BQ_F_CUSTOM = 5,   /* user expression compiled into the same fused kernel (see below);
                          codes BQ_F_CUSTOM + 1 .. BQ_F_CUSTOM + BQ_MAX_CUSTOM name registered ones */
    BQ_F_COUNT = 6
};
#define BQ_MAX_CUSTOM 64

/* Quadrature rule.  The reference only has LEFT (riemann.cpp:36-39). */
enum bq_rule { BQ_RULE_LEFT = 0, BQ_RULE_RIGHT = 1, BQ_RULE_MID = 2, BQ_RULE_TRAP = 3 };

/* Arithmetic.  F64: fp64 evaluation, dd-compensated combines.  F32C: fp32
 * evaluation; per-thread accumulation is compensated — trig integrands in ROTATE
 * mode: fp32 FMA runs of 32 rows over error-diffused row values, promoted to fp64,
 * with one fp64 closing term per part (on a fine grid the partial sums are exact:
 * the result agrees with F64 to ~1e-15); everything else:
 * pairwise + Kahan — and every cross-thread combine is dd like in F64. */
enum bq_dtype { BQ_F64 = 0, BQ_F32C = 1 };

/* How f is evaluated.  DIRECT: x_i formed and f(x_i) called per point (works
 * for every integrand).  ROTATE: for trig integrands, sin/cos of each point
 * are produced by an exact-angle-sum rotation of an accurately computed tile
 * base (3 fp64 ops per point instead of a full sin()).  AUTO picks ROTATE
 * when the integrand supports it and n is large enough to amortise set-up. */
enum bq_eval { BQ_EVAL_AUTO = 0, BQ_EVAL_DIRECT = 1, BQ_EVAL_ROTATE = 2 };

enum bq_status {
    BQ_OK = 0,
    BQ_ERR_INVALID = 1,      /* bad argument */
    BQ_ERR_CUDA = 2,         /* a CUDA runtime call failed; see bq_last_error() */
    BQ_ERR_TIMEOUT = 3,      /* a peer never arrived at the in-kernel barrier */
    BQ_ERR_UNSUPPORTED = 4,
    BQ_ERR_RANGE = 5         /* integrand evaluated outside its table */
};

/* ---- job / result ------------------------------------------------------ */

#define BQ_SEGMENTS 8          /* fixed top-level split of the index space     */
#define BQ_MAX_WORLD 8

typedef struct bq_job {
    int32_t integrand;         /* enum bq_integrand */
    int32_t rule;              /* enum bq_rule */
    int32_t dtype;             /* enum bq_dtype */
    int32_t eval;              /* enum bq_eval */
    double a, b;               /* interval */
    uint64_t n;                /* number of sub-intervals, 1 .. 2^40 */
    /* integrand parameters; 0 means "reference default":
     *  ACC/VEL/DIS: p[0]=tscale p[1]=ascale p[2]=vscale (riemann.cpp:7-9)
     *  PROFILE    : p[0]=table spacing in x units (1.0 s, ex4vel.h:1-5) */
    double p[4];
} bq_job;

typedef struct bq_result {
    double value;              /* the integral */
    double value_lo;           /* low word of the dd result (value+value_lo) */
    double raw_sum_hi;         /* sum of f(x_i) before multiplying by h */
    double raw_sum_lo;
    double seg_hi[BQ_SEGMENTS];/* the 8 fixed segment sums (debug / parity) */
    double seg_lo[BQ_SEGMENTS];
    uint64_t t_start_ns;       /* %globaltimer at kernel entry (first CTA)    */
    uint64_t t_end_ns;         /* %globaltimer when the result was written    */
    uint64_t n;                /* echo */
    int32_t status;            /* enum bq_status, written by the kernel       */
    int32_t eval_used;         /* enum bq_eval actually used                  */
    int32_t world;             /* ranks that contributed                      */
    int32_t rank;
    float device_ms;           /* cudaEvent time of the launch (sync calls)   */
    int32_t launches;          /* kernels launched for this call (always 1)   */
    uint64_t seq;              /* call number of the context; in the host copy of the block it is
                                  the LAST word the kernel writes (completion flag)           */
    uint64_t reserved;
    /* Where the kernel's fixed per-call time goes: nanoseconds since t_start_ns, taken from
     * %globaltimer by the CTA that finishes the call (enum bq_stamp).  A stamp the call did not
     * reach (e.g. no cross-GPU stage in a world of 1) repeats the previous one. */
    uint32_t stamp_ns[8];
} bq_result;

enum bq_stamp {
    BQ_STAMP_PROLOGUE = 0,   /* finishing CTA: column constants / table ready, tile loop starts */
    BQ_STAMP_TILES = 1,      /* finishing CTA: its last tile summed */
    BQ_STAMP_TICKET = 2,     /* it drew the last ticket: every chunk sum of this GPU is stored */
    BQ_STAMP_FOLD = 3,       /* segment sums folded */
    BQ_STAMP_PUBLISHED = 4,  /* segment pairs stored into every peer (world > 1) */
    BQ_STAMP_BARRIER = 5,    /* every peer's pairs have arrived */
    BQ_STAMP_RESULT = 6,     /* result block assembled (== t_end_ns - t_start_ns) */
    BQ_STAMP_SPINS = 7       /* not a time: polls of the in-kernel barrier loop */
};

/* ---- cumulative integration (scan) -------------------------------------
 * Capability of 4main.c:76-225: sample f on a uniform grid, running sum
 * P[i] = sum_{k<=i} f(x_k) ("default_sum"), and running sum of that,
 * Q[i] = sum_{k<=i} P[k] ("default_sum_of_sums"), both over the whole grid
 * even when the grid is split over several GPUs. */
typedef struct bq_scan_job {
    int32_t integrand;
    int32_t replicate;         /* 1: every rank ends with the whole table (MPI_Bcast, 4main.c:157,221) */
    int32_t want_samples;      /* 1: also store f(x_i) ("InterpProfile", 4main.c:82-86) */
    int32_t reserved;
    double x0, dx;             /* x_i = x0 + i*dx */
    uint64_t n;                /* samples */
    double p[4];
} bq_scan_job;

typedef struct bq_scan_result {
    double p_last;             /* P[n-1] */
    double q_last;             /* Q[n-1] */
    double p_second_last;      /* P[n-2] — the element 4main.c:241 prints */
    uint64_t slice_begin;      /* this rank's slice of the tables */
    uint64_t slice_end;
    uint64_t t_start_ns, t_end_ns;
    int32_t status;
    int32_t world;
    int32_t rank;
    float device_ms;
    uint64_t reserved[2];
} bq_scan_result;

/* ---- custom integrands --------------------------------------------------
 * The reference hard-codes its integrand (riemann.cpp:37, cintegrate.cu:68).
 * Here a C++ expression in x (and parameters p0..p3) can be compiled into the
 * SAME fused kernel — evaluation, block/cluster reduction, cross-GPU combine —
 * by a small translation unit that instantiates the kernel template for it
 * (b200quad.models.custom(), models/jit.py, generates, builds and loads it).  That unit hands
 * the library two entry points; jobs with integrand = BQ_F_CUSTOM then use them.
 *   launch(plan, device_state, launch_config, stream) -> cudaError_t as int
 *   eval(x, p[4])                                     -> f(x) on the host (CPU backend, oracles)
 * bq_register_custom_integrand returns the integrand CODE to put into bq_job.integrand (or -1 when
 * the BQ_MAX_CUSTOM slots are used up); the pair belongs to the jobs that name it, so any number of
 * expressions can be in use at once, from any thread.  bq_set_custom_integrand is the older
 * single-slot form: it (re)binds code BQ_F_CUSTOM itself, process-wide. */
typedef int (*bq_custom_launch_fn)(const void* plan, const void* device_state, const void* launch_config,
                                   void* stream);
typedef double (*bq_custom_eval_fn)(double x, const double* p);
int bq_register_custom_integrand(bq_custom_launch_fn launch, bq_custom_eval_fn eval);
int bq_set_custom_integrand(bq_custom_launch_fn launch, bq_custom_eval_fn eval);

/* ---- library-level ----------------------------------------------------- */

const char* bq_version(void);
const char* bq_last_error(void);               /* thread-local message */
int bq_cuda_device_count(void);                /* 0 when no driver / no GPU */
const double* bq_default_profile(int* len);    /* the 1801-entry table */
/* Bytes moved host->device per call: the job description (the plan) and, for the fused fp64 ROTATE
 * kernels, the 128 row constants — both ride in the kernel-launch parameter buffer. */
size_t bq_sizeof_params(void);

/* The bounded-argument sin / cos that DIRECT evaluation uses when the plan has verified
 * |angle| <= 1e6 over the whole interval (host build of the same code the kernel runs; replaces the
 * libm sin() call of riemann.cpp:37 / cintegrate.cu:68).  Exposed for the accuracy tests. */
double bq_sin_bounded(double x);
double bq_cos_bounded(double x);

/* Closed-form / exactly-rounded oracles used by the tests and the CLIs. */
double bq_oracle_sin_sum(double a, double b, uint64_t n, int rule);

/* ---- CPU backend (same summation tree as the kernels) ------------------ */

/* Whole job on the host with `threads` worker threads. */
int bq_cpu_integrate(const bq_job* job, int threads, bq_result* out);
/* Only the segments owned by `rank` of `world`; seg_hi/seg_lo of the others
 * are left 0.  Used by the gloo / multi-process CPU path. */
int bq_cpu_integrate_segments(const bq_job* job, int rank, int world, int threads, bq_result* out);
/* Fold 8 segment pairs (in order) + rule epilogue -> value.  The same code the
 * kernel's last stage runs. */
int bq_fold_segments(const bq_job* job, const double* seg_hi, const double* seg_lo, bq_result* out);
int bq_cpu_cumulative(const bq_scan_job* job, int threads, double* P, double* Q, double* samples,
                      bq_scan_result* out);

/* One rank of a multi-process CPU world (gloo layout).  Phase 1: totals[4*s .. 4*s+3] = (P_hi, P_lo,
 * Q_hi, Q_lo) of every segment s this rank owns (rank r of world W owns 8/W consecutive segments);
 * the caller gathers the 8 totals from all ranks (the only communication: 32 doubles, where the
 * reference gathers whole slices to its root, 4main.c:141-155); phase 2 writes this rank's slice
 * [out->slice_begin, out->slice_end) of the tables into buffers that start at the slice. */
int bq_cpu_scan_totals(const bq_scan_job* job, int rank, int world, int threads, double* totals /* 32 */);
int bq_cpu_scan_finish(const bq_scan_job* job, int rank, int world, int threads, const double* totals,
                       double* P_slice, double* Q_slice, double* samples_slice, bq_scan_result* out);

/* ---- CUDA contexts ----------------------------------------------------- */

typedef struct bq_ctx bq_ctx;

/* One context per (process, device).  Allocates: partials, ticket, result
 * blocks (device + pinned mapped host), comm block (peer-visible). */
int bq_ctx_create(int device, bq_ctx** out);
void bq_ctx_destroy(bq_ctx* ctx);
int bq_ctx_device(const bq_ctx* ctx);
/* Stream used by the synchronous calls (a cudaStream_t). */
void* bq_ctx_stream(const bq_ctx* ctx);
/* Replace the PROFILE table (len <= 4096 entries). */
int bq_ctx_set_profile(bq_ctx* ctx, const double* table, int len);
/* Launch configuration knobs (0 = default): CTAs per SM, cluster size (1|2),
 * rows per tile.  For tuning / tests; results do not depend on the first two. */
int bq_ctx_configure(bq_ctx* ctx, int ctas_per_sm, int cluster, int rows_per_tile);
/* In-kernel barrier time-out in milliseconds (default 2000). */
int bq_ctx_set_timeout_ms(bq_ctx* ctx, int ms);

/* Synchronous: ONE kernel launch (the job rides in the launch parameters, the
 * kernel writes the result block into pinned host memory), bracketed by two
 * events for out->device_ms. */
int bq_integrate(bq_ctx* ctx, const bq_job* job, bq_result* out);
/* Lowest-latency synchronous call: no events, and instead of a stream
 * synchronise the host watches the completion word (bq_result.seq) the kernel
 * writes last into the pinned result block.  out may be NULL; *value gets the
 * integral.  This is what the Python integrate() calls. */
int bq_integrate_value(bq_ctx* ctx, const bq_job* job, double* value, bq_result* out);
/* Asynchronous on `stream` (cudaStream_t; NULL = ctx stream): enqueues the one
 * kernel.  bq_result_fetch() synchronises and copies the result out. */
int bq_integrate_async(bq_ctx* ctx, const bq_job* job, void* stream);
int bq_result_fetch(bq_ctx* ctx, bq_result* out);
/* Device-resident variant (what the torch op calls): exactly one kernel, which
 * stores the integral at d_value (device memory, 8-byte aligned) and the full
 * block in the context's device result block; nothing is written to the host,
 * so nothing at the end of the kernel waits for PCIe.  Capturable in a CUDA graph. */
int bq_integrate_device_async(bq_ctx* ctx, const bq_job* job, double* d_value, void* stream);
/* Synchronises and copies the DEVICE result block of the most recent launch out. */
int bq_device_result_fetch(bq_ctx* ctx, bq_result* out);
/* Device pointer of the result block of the most recent launch. */
const bq_result* bq_ctx_device_result(const bq_ctx* ctx);

int bq_cumulative(bq_ctx* ctx, const bq_scan_job* job, double* dP, double* dQ, double* dSamples,
                  bq_scan_result* out);
int bq_cumulative_async(bq_ctx* ctx, const bq_scan_job* job, double* dP, double* dQ,
                        double* dSamples, void* stream);
int bq_scan_result_fetch(bq_ctx* ctx, bq_scan_result* out);
/* The scan's descriptor scratch grows with n on first use (cudaMalloc, which synchronises and
 * cannot be captured): size it for grids of up to n samples ahead of a latency-critical call or
 * a stream capture. */
int bq_ctx_reserve_scan(bq_ctx* ctx, uint64_t n);

/* Cumulative integration that allocates nothing (no counterpart in the
 * reference, which cudaMallocs and frees on every run: cintegrate.cu:116-118,
 * 143-145).  For callers that cannot allocate device memory: real-time loops
 * where cudaMalloc/cudaFree (device-synchronising, slow) are banned, plug-ins
 * inside a host application that owns all device memory and only lends fixed
 * buffers, processes so close to the memory limit that an allocation may fail.
 * Same fused kernel and same results as bq_cumulative on one GPU, but its
 * scratch is a fixed static device array compiled into the library (about
 * 3.3 MB per GPU, held until the process exits), so:
 *   - n must not exceed BQ_NOALLOC_MAX_SAMPLES (larger n is rejected with
 *     BQ_ERR_INVALID and a message; use bq_cumulative instead);
 *   - the first call on a device does that device's one-time set-up
 *     (synchronous runtime calls): make it before any stream capture begins;
 *   - every call on a device shares the array: calls that overlap on the GPU
 *     race, so issue them on one stream (or otherwise serialise them);
 *   - single GPU only; job->replicate and job->want_samples are ignored;
 *   - the PROFILE integrand is always the library's default table (ex4vel.h): a table set with
 *     bq_ctx_set_profile belongs to a context, and this entry point has none.
 * The caller passes the inputs (job) and the outputs: P and Q (device memory,
 * n doubles each, 32-byte aligned) and, optionally, `out` (device or pinned
 * host memory) which receives the result block.  Work is enqueued on `stream`
 * (a cudaStream_t, NULL = default stream) and NOT waited for: synchronise the
 * stream before reading P, Q or *out. */
#define BQ_NOALLOC_MAX_SAMPLES 67108864ull     /* 32768 tiles of 2048 samples */
uint64_t bq_noalloc_max_samples(void);
int bq_cumulative_noalloc(int device, const bq_scan_job* job, double* dP, double* dQ,
                          bq_scan_result* out, void* stream);

/* ---- multi-GPU: peer wiring --------------------------------------------
 * Replaces MPI_Send/Recv (riemann.cpp:76,83), MPI_Reduce, MPI_Barrier and
 * MPI_Bcast (4main.c:134,137,157,200,221): the fused kernels publish their
 * partials into every peer's comm block with NVLink stores and meet at an
 * in-kernel epoch barrier. */
#define BQ_IPC_HANDLE_BYTES 64
/* Multi-process: export this context's comm block ... */
int bq_ctx_export_comm(bq_ctx* ctx, unsigned char handle[BQ_IPC_HANDLE_BYTES]);
/* ... and attach the other ranks' (handles[r] for r != rank; own slot ignored). */
int bq_ctx_attach_peers(bq_ctx* ctx, int rank, int world,
                        const unsigned char* handles /* world * 64 bytes */);
/* Alternative wiring for callers that already hold symmetric memory (e.g. a
 * torch.distributed._symmetric_memory rendezvous): comm_ptrs[r] is rank r's block (>= 1 KiB,
 * 256-byte aligned, `bytes` long) as mapped in THIS process, mc_ptr the NVLS multicast address of the
 * same blocks or NULL.  The context uses comm_ptrs[rank] as its own block (and clears it): every
 * rank must have returned from this call before the first collective call (host barrier). */
int bq_ctx_adopt_comm(bq_ctx* ctx, int rank, int world, void* const* comm_ptrs, void* mc_ptr, size_t bytes);
int bq_ctx_set_nvls(bq_ctx* ctx, int on);
/* Multicast addresses of the replicated output tables of the next scans (NULL = peer stores). */
int bq_ctx_set_scan_multicast(bq_ctx* ctx, double* mcP, double* mcQ);
/* Zero this context's barrier state.  Every rank of a world must call it, and
 * the caller must make sure all have done so (host barrier) before the next
 * collective call; needed after a BQ_ERR_TIMEOUT. */
int bq_ctx_reset_comm(bq_ctx* ctx);
/* Same for an output table the scan replicates into (per call buffer). */
int bq_ipc_export(void* dptr, unsigned char handle[BQ_IPC_HANDLE_BYTES]);
/* Export of a pointer that may lie INSIDE a larger allocation (e.g. a tensor handed out by a
 * caching allocator): *offset is dptr's distance from the base of its allocation, which is what
 * bq_ipc_open returns on the importing side — the peer's view of dptr is base + *offset. */
int bq_ipc_export_at(void* dptr, unsigned char handle[BQ_IPC_HANDLE_BYTES], uint64_t* offset);
int bq_ipc_open(const unsigned char handle[BQ_IPC_HANDLE_BYTES], void** dptr);
int bq_ipc_close(void* dptr);
int bq_ctx_set_scan_peers(bq_ctx* ctx, double* const* peerP, double* const* peerQ);

/* Single-process world: one context per listed device, peer access enabled. */
typedef struct bq_world bq_world;
int bq_world_create(int ndev, const int* devices, bq_world** out);
void bq_world_destroy(bq_world* w);
int bq_world_size(const bq_world* w);
bq_ctx* bq_world_ctx(bq_world* w, int rank);
/* NVLS (NVSwitch multicast): a world whose comm blocks could be bound to a multicast object can
 * publish with one multimem.st per word, fanned out by the switch, instead of one store per peer.
 * Off by default (BQ_NVLS=1 in the environment turns it on at creation); same results either way. */
int bq_world_has_multicast(const bq_world* w);
int bq_world_set_nvls(bq_world* w, int on);
/* Output tables of n doubles per rank in symmetric memory, bound to a multicast object when the
 * fabric has one: a replicating bq_world_cumulative on them does its fused all-gather with
 * multimem.st (when NVLS is on) instead of world-1 peer stores per datum.  P[r] / Q[r] receive
 * rank r's table.  Freed by bq_world_free_tables or with the world. */
int bq_world_alloc_tables(bq_world* w, uint64_t n, double** P, double** Q);
void bq_world_free_tables(bq_world* w);
double* bq_world_table_multicast(bq_world* w, const double* rank0_ptr);
/* Launch on every device, wait for all; out[r] for each rank (may be NULL). */
int bq_world_integrate(bq_world* w, const bq_job* job, bq_result* out);
int bq_world_cumulative(bq_world* w, const bq_scan_job* job, double* const* dP, double* const* dQ,
                        bq_scan_result* out);

/* Process-wide defaults, created on first use and never destroyed: the context of a device, the
 * single-process world of a size, and (set by the Python DistWorld) the context that is this
 * process's rank of the torch.distributed world.  The Python layer and the torch ops share them. */
bq_ctx* bq_default_ctx(int device);
bq_world* bq_default_world(int size);
int bq_set_dist_ctx(bq_ctx* ctx);
bq_ctx* bq_dist_ctx(void);
int bq_ctx_world(const bq_ctx* ctx);
int bq_ctx_rank(const bq_ctx* ctx);

/* Baseline comparator pieces (NOT the product path): the same kernel with the
 * cross-GPU stage disabled leaves its 8/world segment pairs in a device
 * buffer so an external all-reduce (NCCL) can finish the job. */
int bq_integrate_local_async(bq_ctx* ctx, const bq_job* job, int rank, int world,
                             double* dSegPairs /* 2*BQ_SEGMENTS doubles */, void* stream);

/* Device memory for C callers that do not link the CUDA runtime themselves. */
int bq_device_alloc(int device, size_t bytes, void** out);
int bq_device_free(int device, void* p);
int bq_device_read(int device, void* host_dst, const void* dev_src, size_t bytes);

/* Tracing: when enabled every CTA of the quad kernel records (start ns, end ns,
 * SM id, tiles processed) from %globaltimer; read back after a call.  The
 * reference's only instrumentation is one wall-clock pair around main()
 * (riemann.cpp:51,91; cintegrate.cu:104,139). */
int bq_ctx_trace(bq_ctx* ctx, int enable);
int bq_ctx_trace_read(bq_ctx* ctx, uint64_t* out /* 4 per CTA */, int max_ctas);

/* Fault injection for the time-out path: the next launch on this context
 * skips its publish step (its peers must report BQ_ERR_TIMEOUT). */
int bq_ctx_inject_fault(bq_ctx* ctx, int kind);

#ifdef __cplusplus
}
#endif
#endif /* B200QUAD_H_ */
