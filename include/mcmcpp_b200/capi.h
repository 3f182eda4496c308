/*
 * include/mcmcpp_b200/capi.h -- C-ABI of libmcmcpp_b200.so, the B200 (sm_100a) runtime behind
 * MCMCpp's ensemble-update hot path.
 *
 * Plain C: opaque handle, plain pointers and sizes, int status returns (0 = ok).  No torch,
 * no C++ types.  Each entry point names the reference interface it replaces (paths relative to
 * the reference checkout, jmatta1/MCMCpp).  The C++ drop-in templates in
 * include/mcmcpp_b200/EnsembleSampler.h and the Python mirror in mcmcpp_b200/sampler.py are thin
 * callers of exactly these functions; INTEGRATION.md shows the binding a maintainer adds.
 *
 * There is no CPU fallback: every call fails with MCMCPP_ERR_CUDA when no sm_100-class device
 * is usable.
 */
#ifndef MCMCPP_B200_CAPI_H
#define MCMCPP_B200_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcmcpp_sampler mcmcpp_sampler;

enum {
    MCMCPP_OK = 0,
    MCMCPP_CHAIN_FULL = 1,      /* runMCMC would return false (EnsembleSampler.h:293,306) */
    MCMCPP_ERR_ARG = -1,        /* the reference asserts (EnsembleSampler.h:207-208,327,335-336) */
    MCMCPP_ERR_CUDA = -2,
    MCMCPP_ERR_NCCL = -3,
    MCMCPP_ERR_STATE = -4,      /* call order (e.g. run before set_initial) */
    MCMCPP_ERR_UNSUPPORTED = -5
};

/* RNG modes.  SEQUENTIAL reproduces EnsembleSampler's single pcg64 stream (seed, stream 0)
 * consumed in walker order (EnsembleSampler.h:217,345-354): walker slot i of half h at step t
 * uses draws 3(tW + hK + i)+1..+3, reached by jump-ahead, so results are bit-identical to the
 * sequential reference and independent of the GPU count.  PER_WALKER gives walker w its own
 * stream pcg64(seed, w) (the per-thread streams of RedBlkUpdater.h:86 taken to one walker per
 * "thread"); its oracle is one reference StretchMove per walker. */
enum { MCMCPP_RNG_SEQUENTIAL = 0, MCMCPP_RNG_PER_WALKER = 1,
       /* the fast engine: walker w owns pcg32(seed, w) (setseq XSH-RR 64/32, pcg_random.hpp:808-836,1568) and draws
        * what the reference's facade (Utility/MultiSampler.h:60,86,99) makes of a 32-bit engine: bounded_rand in 32-bit
        * arithmetic, generate_canonical from two outputs.  Stretch move only.  Oracle: oracle_stretch_run_per_walker32,
        * pinned on pcg-cpp's check-pcg32.out and on the vendored pcg32 + libstdc++ distributions (oracle/ref_driver.cpp). */
       MCMCPP_RNG_PER_WALKER32 = 2 };

/* How one ensemble is spread over `world` GPUs (one process per GPU):
 *   NONE       world == 1
 *   ALLGATHER  walker-index shards; after each half-step the updated half is all-gathered with
 *              NCCL so every rank holds the full ensemble (BASELINE.json north_star)
 *   PEER       walker-index shards; partner rows are read straight from the owning GPU's HBM
 *              over NVLink (peer-mapped memory) inside the update kernel; one device-side
 *              barrier per half-step (epoch flags behind every rank's position array, stored over
 *              NVLink by a one-CTA kernel; no NCCL call), no bulk exchange
 *   DATA       the Calculator's DATA is sharded, the ensemble is replicated: every rank makes the same draws
 *              and proposals for all W walkers, evaluates its partial log-likelihoods (its rows of the data
 *              matrix, or its slice of a separable sum), the K partials of a colour are summed over ranks with
 *              one ncclAllReduce (8 K bytes) and every rank takes the same decisions.  The
 *              launcher calls args->allreduce_sum between its partial and decision kernels.
 *   DIM        the PARAMETER dimension is sharded, the walkers are replicated: rank r holds columns
 *              [dim_begin, dim_begin + D_local) of every walker row (even-sized slices, see
 *              mcmcpp_dim_slice), makes the same draws as everyone, and a separable Calculator's partial sums are
 *              exchanged by the update kernel itself through peer-mapped tables (8-byte NVLink stores polled by
 *              their one reader; no flag, no fence, no NCCL), summed in rank
 *              order (include/mcmcpp_b200/dim_shard_kernels.cuh).  Traffic per rank is 1/world of the
 *              single-GPU kernel's and the exchange does not grow with D.  Host-facing arrays keep the full
 *              [W][num_params] shape; a rank reads/writes only its columns of them. */
enum { MCMCPP_SHARD_NONE = 0, MCMCPP_SHARD_ALLGATHER = 1, MCMCPP_SHARD_PEER = 2, MCMCPP_SHARD_DATA = 3, MCMCPP_SHARD_DIM = 4 };

/* built-in Calculators (include/mcmcpp_b200/targets.h); params layout as in oracle/oracle.c */
enum {
    MCMCPP_TARGET_SKEWED_GAUSS_2D = 0, /* params {eps}             test/sequential/SkewedGaussian/Common/SkewedGaussian.h:52-57 */
    MCMCPP_TARGET_ROSENBROCK_2D = 1,   /* params {}                */
    MCMCPP_TARGET_DIAG_GAUSS = 2,      /* params iv[D]             */
    MCMCPP_TARGET_DENSE_GAUSS = 3,     /* params P[D*D] row-major  */
    MCMCPP_TARGET_LINREG = 4           /* params {M, priorVar, X[M*D] row-major, y[M]} */
};

typedef struct {
    int32_t num_walkers;       /* numWalker: even, > 2*num_params  (EnsembleSampler.h:66,207-208) */
    int32_t num_params;        /* numParameter */
    int64_t seed;              /* randSeed -> setPrng(seed, stream)  (EnsembleSampler.h:217) */
    int32_t rng_mode;          /* MCMCPP_RNG_* */
    int32_t alpha_num;         /* GwDistribution<ParamType, AlphaNum, AlphaDenom>  (Utility/GwDistribution.h:45) */
    int32_t alpha_denom;
    int32_t device;            /* CUDA device ordinal */
    int32_t rank;              /* this process's shard */
    int32_t world;             /* number of shards (GPUs) */
    int32_t shard_mode;        /* MCMCPP_SHARD_* */
    int32_t trace_updates;     /* >0: record (j, accept, z, e, newp, diff) for the first n updates (tests) */
    uint64_t max_chain_bytes;  /* maxChainSizeBytes (EnsembleSampler.h:67); 0 = keep no chain */
    int32_t mover_state_doubles; /* doubles of per-walker mover state the runtime keeps between updates (0 for the stretch move;
                                  * 2 for WalkMove: its normal distribution's saved value + flag); zeroed at creation */
    int32_t chain_block_steps;   /* 0: the chain holds floor(max_chain_bytes / (8 W D)) rows.  n > 0: the reference's granularity --
                                  * whole blocks of n steps, max(1, floor(max_chain_bytes / (8 W D n))) of them
                                  * (Chain/Chain.h:178, ChainBlock.h:31 BlockSize = 10000; the first block always exists, Chain.h:181):
                                  * runMCMC returns false at the step the reference returns false at */
    uint64_t device_chain_bytes; /* > 0: stored steps stay in HBM (up to this many bytes of rows) and cross PCIe only when the host
                                  * asks for them (mcmcpp_sampler_chain) or the device block is full (then they spill to the pinned
                                  * host chain); sliceAndBurnChain compacts them on the device; covariance and autocorrelation times
                                  * can be taken from them in place (mcmcpp_sampler_chain_covariance / _autocorr_times).  With
                                  * max_chain_bytes == 0 the device block IS the chain (runMCMC returns false when it is full).
                                  * 0: every stored row is streamed to the host as it is made (2-row device ring). */
} mcmcpp_config;

/* One row of the optional trace, in the sequential update order (step, half, slot). */
typedef struct {
    int32_t j;
    int32_t accept;
    double z, e, newp, diff;
} mcmcpp_trace;

/*
 * Arguments of one half-step launch.  A "launcher" is a host function that enqueues the
 * update kernel for the owned slots of one colour on `stream`; the runtime owns every buffer.
 * Launchers for user Calculators are instantiated in the user's CUDA translation unit by
 * include/mcmcpp_b200/stretch_kernels.cuh (MCMC::Device::HalfStepLauncher<Calc>::launch);
 * launchers for the built-in targets live inside the library.
 *
 * Layout: walker-major rows.  pos[w*D + d], walkers 0..K-1 red, K..W-1 black
 * (EnsembleSampler.h:225-226) -- the reference's own chain-row layout (ChainBlock.h:125-131),
 * so a stored step is a plain copy of the state.
 */
/* DATA sharding: sums dev_buf[0..count) over all ranks in place, ordered on cuda_stream; every rank receives the
 * same bits. */
typedef int (*mcmcpp_allreduce_fn)(void* ctx, double* dev_buf, size_t count, void* cuda_stream);

typedef struct {
    double* pos;               /* [W*D] local full-ensemble array (own rows always valid; all rows valid unless PEER) */
    double* logp;              /* [W]   auxData of owned walkers */
    uint64_t* rng_state;       /* [W*2] pcg64 state (lo,hi) per walker, positioned at its next draw */
    double* chain_row;         /* device chain slot [W*D] for this step, or NULL when the step is not stored */
    unsigned long long* accept_slots; /* [grid] per-CTA accepted-proposal partial sums */
    int32_t* flags;            /* [0]: count of bounded_rand redraws seen (sequential-stream cadence broken);
                                * [1]: a cross-GPU wait timed out (mcmcpp_sampler_run reports it) */
    mcmcpp_trace* trace;       /* trace row of slot 0 of this half-step, or NULL */
    const double* const* peer_pos; /* PEER mode: [world] device pointers to every rank's pos array, else NULL */
    int32_t W, D, K;
    int32_t half;              /* 0 red, 1 black */
    int32_t slot_begin;        /* first owned slot in [0,K) */
    int32_t slot_count;        /* owned slots */
    int32_t slots_per_rank;    /* K / world (PEER owner lookup) */
    int32_t rng_mode;
    uint64_t bound_threshold;  /* (2^64 - K) mod K   (pcg_extras.hpp:517); with the 32-bit engine (PER_WALKER32): (2^32 - K) mod K */
    uint64_t bound_recip;      /* floor(2^64 / K) */
    uint64_t skip_mult[2];     /* SEQUENTIAL: affine jump by 3W-3 draws applied after a walker's 3 draws */
    uint64_t skip_plus[2];
    double gw_term1;           /* GwDistribution constants */
    double gw_inv_sqrt_alpha;
    int32_t max_grid;          /* size of accept_slots */
    int32_t trace_rows;        /* rows available at `trace` (slots >= trace_rows are not recorded) */
    mcmcpp_allreduce_fn allreduce_sum; /* DATA sharding with world > 1, else NULL */
    void* allreduce_ctx;
    unsigned long long* launch_count;  /* host counter: a launcher that enqueues n kernels adds n-1 (the runtime adds 1) */
    /* DIM sharding (D above is then this rank's slice length) */
    int32_t D_total;           /* numParams of the whole row: log(z) * (D_total - 1)  (StretchMove.h:110) */
    int32_t dim_begin;         /* global index of local column 0 */
    int32_t rank, world;
    double* const* dim_partial;/* [world] every rank's partial table [2 colours][world][K] (own entry local, others peer-mapped);
                                * a slot is all-ones (empty) or a partial sum -- see dim_shard_kernels.cuh */
    double* dim_z;             /* [2K] draw scratch (a launcher may double-buffer by colour): draws of the colour in flight, kept between kernels (all shard modes) */
    double* dim_e;
    int32_t* dim_j;
    double* dim_s;             /* [2K] log(z) (D_total - 1) of the draws (DIM), second half free for a launcher's own use */
    int32_t sm_count;          /* multiprocessors of the device (persistent grids) */
    double* mover_state;       /* [W * mover_state_doubles] per-walker mover state, or NULL */
    void* side_stream;         /* a second cudaStream_t of the sampler and two cudaEvent_t (timing disabled) for launchers that want to */
    void* side_event[2];       /* run part of a half-step beside the rest; the built-in launchers do not use them (DIM: measured slower) */
} mcmcpp_halfstep_args;

typedef int (*mcmcpp_halfstep_fn)(const mcmcpp_halfstep_args* args, const void* calculator, void* cuda_stream);

/* Small ensembles (at most 1024 walkers per colour, at most 8 parameters): the loop of EnsembleSampler::runMCMC
 * (EnsembleSampler.h:284-310) -- `stored_steps` x (`interval` - 1 unsaved ensemble steps + 1 saved) -- in ONE launch of
 * one CTA, walker state in shared memory, a CTA barrier where the reference has its colour boundary.  At 32..2048
 * walkers a half-step is a microsecond of arithmetic behind 3 us of launch latency; same arithmetic, same results. */
typedef struct mcmcpp_fused_args {
    mcmcpp_halfstep_args base;  /* as for a half-step; half, chain_row and trace are not used */
    int32_t stored_steps;       /* stored steps to make */
    int32_t interval;           /* ensemble steps per stored step */
    double* chain_rows;         /* [stored_steps][W][D] device rows to fill, or NULL (chain off / full) */
} mcmcpp_fused_args;
typedef int (*mcmcpp_fused_fn)(const mcmcpp_fused_args* args, const void* calculator, void* cuda_stream);

/* All ranks of one DIM-sharded ensemble held by ONE process on ONE device (mcmcpp_sampler_run_group): one colour of every rank --
 * every rank's draws and summing kernel on `cuda_stream`, then every rank's finishing kernel on its side stream
 * (include/mcmcpp_b200/dim_shard_kernels.cuh: finishing kernels wait for summing kernels only, which wait for nobody). */
typedef int (*mcmcpp_group_fn)(const mcmcpp_halfstep_args* args_per_rank, const void* const* calculators, int count, void* cuda_stream);

/* ---- life cycle: EnsembleSampler ctor/dtor (EnsembleSampler.h:199-218,72) -------------------- */
/* `calculator` is copied by value (calc_bytes); it must be trivially copyable and any pointers in
 * it must be device pointers (the Calculator "view" convention of targets.h). */
int mcmcpp_sampler_create(const mcmcpp_config* cfg, mcmcpp_halfstep_fn launcher, const void* calculator,
                          size_t calc_bytes, mcmcpp_sampler** out);
/* Optional: the fused small-ensemble launcher for the same mover and Calculator (NULL removes it).  The runtime uses it
 * when the ensemble fits (see mcmcpp_fused_args), one GPU, no trace, no step action and no on-the-fly covariance;
 * MCMCPP_B200_FUSED=0 in the environment keeps the launch-per-colour path. */
int mcmcpp_sampler_set_fused_launcher(mcmcpp_sampler* s, mcmcpp_fused_fn launcher);
/* Built-in Calculator: `params` (host) is uploaded and viewed by the functor. */
int mcmcpp_sampler_create_builtin(const mcmcpp_config* cfg, int target_kind, const double* params, size_t num_params_values,
                                  mcmcpp_sampler** out);
/* Movers.  STRETCH = Mover::StretchMove (Movers/StretchMove.h); DIFF_EVOLUTION = Mover::DifferentialEvolution
 * (Movers/DifferentialEvolution.h:113-142), mover_params {gamma (<= 0: 2.38/sqrt(2D), :58), rand (<= 0: 1e-4, :83)};
 * WALK = Mover::WalkMove (Movers/WalkMove.h:95-186), mover_params {numSamples (1..64, <= walkers per colour)}.
 * Their draw counts per update vary, so they need rng_mode PER_WALKER; walker sharding only.  User Calculators:
 * MCMC::Device::DiffEvoLauncher<Calc> / WalkMoveLauncher<Calc> (include/mcmcpp_b200/diff_evo_kernels.cuh,
 * walk_move_kernels.cuh) through mcmcpp_sampler_create (WalkMove: cfg.mover_state_doubles = 2). */
enum { MCMCPP_MOVER_STRETCH = 0, MCMCPP_MOVER_DIFF_EVOLUTION = 1, MCMCPP_MOVER_WALK = 2 };
int mcmcpp_sampler_create_builtin_mover(const mcmcpp_config* cfg, int mover_kind, const double* mover_params, size_t num_mover_params,
                                        int target_kind, const double* params, size_t num_params_values, mcmcpp_sampler** out);
int mcmcpp_sampler_destroy(mcmcpp_sampler* s);

/* ---- setInitialWalkerPos (EnsembleSampler.h:220-230): AoS [walker][param] + aux[walker];
 *      counts as one accepted step per walker and stores chain row 0 (Walker.h:76,162-170) -------- */
int mcmcpp_sampler_set_initial(mcmcpp_sampler* s, const double* positions_host, const double* aux_host);
/* same with device-resident inputs (no host copy) */
int mcmcpp_sampler_set_initial_device(mcmcpp_sampler* s, const double* positions_dev, const double* aux_dev);

/* DIM sharding: the same with this rank's columns only, contiguous [W][D_local] (mcmcpp_dim_slice) -- each process
 * keeps its shard of the rows on its own host side; in every other mode identical to the full-row calls.  Read side:
 * mcmcpp_sampler_get_state_local. */
int mcmcpp_sampler_set_initial_local(mcmcpp_sampler* s, const double* positions_local_host, const double* aux_host);
int mcmcpp_sampler_get_state_local(mcmcpp_sampler* s, double* positions_local_host, double* aux_host);

/* ---- setSlicingMode / setSamplingMode (EnsembleSampler.h:324-330, ParallelEnsembleSampler.h:322-330) */
int mcmcpp_sampler_set_slicing(mcmcpp_sampler* s, int use_slicing, int interval);

/* ---- runMCMC (EnsembleSampler.h:284-310): num_steps STORED steps (x interval when slicing).
 *      Returns MCMCPP_CHAIN_FULL where the reference returns false.  Blocking; the chain rows of
 *      the call are on the host when it returns.  device_ms (may be NULL) receives the CUDA-event
 *      time of the update kernels + exchanges on the sampler's stream. */
int mcmcpp_sampler_run(mcmcpp_sampler* s, int num_steps, float* device_ms);

/* ---- getAcceptedSteps / getTotalSteps / getAcceptanceFraction (EnsembleSampler.h:244-282);
 *      ensemble-wide (summed over ranks by the caller when world > 1: values are this rank's) ---- */
int mcmcpp_sampler_get_counts(mcmcpp_sampler* s, unsigned long long* accepted, unsigned long long* total);
int mcmcpp_sampler_get_stored_steps(mcmcpp_sampler* s);                  /* getStoredSteps (EnsembleSampler.h:150) */

/* ---- reset / storeCurrentWalkerPositions / sliceAndBurnChain (EnsembleSampler.h:312-322,232-242,332-339) */
int mcmcpp_sampler_reset(mcmcpp_sampler* s);
int mcmcpp_sampler_store_current(mcmcpp_sampler* s);
int mcmcpp_sampler_slice_and_burn(mcmcpp_sampler* s, int interval, int burn_in);

/* ---- chain read side: contiguous host rows [stored][W][D], the layout ChainStepIterator
 *      dereferences to (ChainStepIterator.h:125).  Valid until reset/slice/destroy/run. ---------- */
int mcmcpp_sampler_chain(mcmcpp_sampler* s, const double** rows_host, long long* stored_steps);
/* Drops the first `rows` host rows (they have been copied elsewhere): the runtime's pinned rows become a bounded staging
 * window in front of a chain the CALLER owns -- the C++ drop-in header uses it to fill the reference's own
 * Chain::Chain / ChainBlock list (Chain/Chain.h:204-236), so the reference's iterators and Analysis classes run on it
 * unchanged.  Counters and walker state are untouched; getStoredSteps goes down by `rows`. */
int mcmcpp_sampler_chain_consume(mcmcpp_sampler* s, long long rows);

/* Device-resident chain (cfg.device_chain_bytes > 0): the rows currently held in HBM -- the LAST *rows stored steps, [rows][W][D_local]
 * contiguous; *first_step is the chain index of the first of them (0 unless older rows have spilled to the host). */
int mcmcpp_sampler_chain_device(mcmcpp_sampler* s, const double** rows_dev, long long* first_step, long long* rows);
/* Analysis over the device-resident chain, nothing crosses PCIe but the results (one GPU or DATA sharding; the whole chain
 * must still be in HBM).  Definitions and argument meaning as mcmcpp_covariance / mcmcpp_autocorr_times below.  With
 * reference_buffer_reuse == 0 and many walkers the autocovariances are averaged in walker groups (one CTA per parameter and
 * group; compensated sums inside a group and across groups): same definition, summation order differs from the sequential
 * walker loop (~1e-13 relative). */
int mcmcpp_sampler_chain_covariance(mcmcpp_sampler* s, int slice_interval, double* mean, double* cov, double* corr, long long* points);
int mcmcpp_sampler_chain_autocorr_times(mcmcpp_sampler* s, int window_scaling, int reference_buffer_reuse, double* iat_out);

/* ---- current walker state (Walker::getCurrState / getCurrAuxData, Walker.h:128,134) ---------- */
int mcmcpp_sampler_get_state(mcmcpp_sampler* s, double* positions_host, double* aux_host);
int mcmcpp_sampler_state_device(mcmcpp_sampler* s, double** positions_dev, double** aux_dev);

/* ---- diagnostics -------------------------------------------------------------------------- */
int mcmcpp_sampler_get_trace(mcmcpp_sampler* s, mcmcpp_trace* out, int max_rows, int* rows);
int mcmcpp_sampler_rng_redraws(mcmcpp_sampler* s, int* redraws);
int mcmcpp_sampler_launch_count(mcmcpp_sampler* s, unsigned long long* kernel_launches);
void* mcmcpp_sampler_stream(mcmcpp_sampler* s);      /* cudaStream_t the kernels are launched on */
const char* mcmcpp_last_error(void);
int mcmcpp_device_count(void);
/* measured FP64 tensor-core (DMMA m8n8k4) peak of `device` in TFLOP/s: the roofline denominator of the dense-likelihood kernels */
int mcmcpp_fp64_tensor_peak(int device, double* tflops);
/* self-check of the kernels' log (device_rng.cuh logOfNormal) against CUDA's log() on `device`: both evaluated on x[0..n), which must
 * be positive, finite and normal; the outputs are host arrays of n doubles.  Test infrastructure for the "same bits" claim. */
int mcmcpp_device_log_check(int device, const double* x, size_t n, double* kernel_log, double* cuda_log);

/* ---- multi-GPU plumbing (one process per GPU) -------------------------------------------------
 *      ALLGATHER / DATA: rank 0 makes an id, the host layer broadcasts the 128 bytes, every rank joins (NCCL).
 *      PEER / DIM: every rank exports a 64-byte IPC handle (PEER: its pos array + barrier flags; DIM: its partial table),
 *      the host layer all-gathers them, every rank imports the others'.  No NCCL communicator is needed: the ranks
 *      may even be processes sharing one GPU (CUDA IPC works there; NCCL does not). */
int mcmcpp_comm_unique_id(void* id128);
int mcmcpp_sampler_comm_init(mcmcpp_sampler* s, const void* id128);
int mcmcpp_sampler_peer_export(mcmcpp_sampler* s, void* handle64);
int mcmcpp_sampler_peer_import(mcmcpp_sampler* s, const void* handles64_all_ranks);
/* The ranks of one DIM-sharded ensemble inside ONE process on ONE device (a one-GPU box, tests): samplers[r] was created with
 * rank r, world `count` (<= 8), shard_mode DIM.  _peer_import_local wires their partial tables together by plain pointers;
 * _run_group is runMCMC for all of them in lockstep -- per colour every rank's summing kernel is launched before any rank's finishing
 * kernel (the only kernels that wait: for summing kernels, which wait for nobody), with exactly the kernels one process per GPU runs.
 * Same results as `count` processes on `count` GPUs.  Built-in Calculators register their group launcher themselves. */
int mcmcpp_sampler_set_group_launcher(mcmcpp_sampler* s, mcmcpp_group_fn launcher);
int mcmcpp_sampler_peer_import_local(mcmcpp_sampler* const* samplers, int count);
int mcmcpp_sampler_run_group(mcmcpp_sampler* const* samplers, int count, int num_steps);

/* DIM sharding: columns [*begin, *begin + *count) of a num_params-wide row held by `rank` (pairs of columns are dealt
 * out so that every slice but possibly the last has even length: 16-byte aligned slices for vector loads). */
int mcmcpp_dim_slice(int num_params, int rank, int world, int* begin, int* count);

/* ---- PostStepAction (EnsembleSampler.h:52-56,356-359; Utility/NoAction.h:34-46) with a DEVICE-resident view of the
 *      step: after the black half-step of every stored step the runtime calls fn(ctx, dev_row, step_index, stream) on
 *      the host, where dev_row is this rank's [W][D_local] state = the row just stored and `stream` the cudaStream_t
 *      the update kernels run on.  Work the action enqueues on `stream` sees exactly that step and delays the next one;
 *      the pointer must not be used after work enqueued later on the stream.  NULL removes the action. */
typedef void (*mcmcpp_step_fn)(void* ctx, const double* dev_row, long long step_index, void* cuda_stream);
int mcmcpp_sampler_set_step_action(mcmcpp_sampler* s, mcmcpp_step_fn fn, void* ctx);

/* ---- checkpoint / resume (no counterpart: the reference's IO tree is empty).  The buffer holds this rank's walker
 *      positions, log-probs, every walker's pcg64 state, the step index and the counters; a sampler created with the
 *      same configuration and restored from it continues bit for bit.  The host chain is not part of it. */
int mcmcpp_sampler_checkpoint_bytes(mcmcpp_sampler* s, size_t* bytes);
int mcmcpp_sampler_checkpoint_save(mcmcpp_sampler* s, void* buffer, size_t bytes);
int mcmcpp_sampler_checkpoint_restore(mcmcpp_sampler* s, const void* buffer, size_t bytes);

/* ---- Analysis::CovarianceMatrix (Analysis/CovarianceMatrix.h:120-215) on the device ------------------------
 *      Same definitions as the reference (n = used steps x walkers, cov = E[xy] - E[x]E[y], Kahan-summed,
 *      correlation normalised by rows then columns); the points are summed in a different order (per-CTA Kahan
 *      pairs, CTA order), so values agree with calculateCovar to ~1e-15 relative on the sums, not bit for bit.
 *      num_params <= 128.  Outputs (host): mean[D], cov[D*D], corr[D*D] row-major; any may be NULL.
 *      cov_begin: zeroStorage + start adding every slice_interval-th step STORED by later mcmcpp_sampler_run calls
 *      (the chain itself may be off); include_current != 0 counts the present walker positions as step 0, as the
 *      chain's first row is (EnsembleSampler.h:228).  One GPU, or DATA sharding (replicated ensemble). */
int mcmcpp_sampler_cov_begin(mcmcpp_sampler* s, int slice_interval, int include_current);
int mcmcpp_sampler_cov_result(mcmcpp_sampler* s, double* mean, double* cov, double* corr, long long* points);
int mcmcpp_sampler_cov_end(mcmcpp_sampler* s);
/* calculateCovar(begin, end, sliceInterval) over a host chain [steps][num_walkers][num_params] */
int mcmcpp_covariance(const double* chain_host, long long steps, int num_walkers, int num_params, int slice_interval, int device,
                      double* mean, double* cov, double* corr, long long* points);

/* ---- covariance of rows already in device memory, WITHOUT any allocation.  No counterpart in the reference: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 *      loops, harnesses that forbid it).  Same kernels, definitions and tolerance tier as mcmcpp_covariance; the per-CTA Kahan
 *      pairs and the totals live in a file-scope `static __device__` array of the library (csrc/analysis.cu; about 2.7 MB, held
 *      by every GPU that loads the library until the process exits) instead of cudaMalloc'ed memory.
 *      rows_dev: DEVICE pointer, [num_points][num_params] doubles on `device`; mean[D], cov[D*D], corr[D*D]: HOST outputs,
 *      any may be NULL; cuda_stream: the cudaStream_t to work on (NULL = the default stream); the call returns after the
 *      stream has drained.  num_params <= MCMCPP_NOALLOC_COV_MAX_PARAMS, else MCMCPP_ERR_UNSUPPORTED with a message.
 *      Calls no cudaMalloc / cudaFree / allocating routine.  Every call on a device shares the one array (cleared at the
 *      start of each call): calls that overlap on the same GPU -- two host threads, two streams -- race; serialise them. */
#define MCMCPP_NOALLOC_COV_MAX_PARAMS 32
int mcmcpp_covariance_device_noalloc(const double* rows_dev, long long num_points, int num_params, int device,
                                     double* mean, double* cov, double* corr, void* cuda_stream);

/* ---- Analysis::AutoCorrCalc::calcAutoCorrTimes + retrieveAutoCorrelationTime (Analysis/AutoCorrCalc.h:77,89,150-203;
 *      Detail/AutoCov.h:131-290) over a host chain [steps][num_walkers][num_params], all walkers: iat_out[num_params].
 *      The reference's own arithmetic in its own order (radix-2 circular FFT autocovariance, its twiddles, its
 *      compensated sums, window = window_scaling, default 4): bit-identical to the CPU restatement.
 *      reference_buffer_reuse != 0 also reproduces transferWalker adding each chain onto the buffer the previous
 *      walker's autocovariance was left in (AutoCorrCalc.h:228-232) -- the reference as written, one CTA, in order;
 *      0 = the clean definition (series = chain), one CTA per parameter.  A negative value = the window never closed,
 *      as in the reference (AutoCorrCalc.h:200-202). */
int mcmcpp_autocorr_times(const double* chain_host, long long steps, int num_walkers, int num_params, int window_scaling,
                          int reference_buffer_reuse, int device, double* iat_out);

#ifdef __cplusplus
}
#endif
#endif
