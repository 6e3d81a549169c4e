/*
 * sim_cuda.cu -- the CUDA simulation pipeline behind include/nbody_sim.h.
 *
 * Replaces the reference's Vulkan simulator (ref src/lib/sim_gpu.c:1-361) and its Vulkan
 * context/allocator (ref src/lib/vulkan_ctx.c) with: cudaSetDevice + cudaMalloc'd SoA buffers,
 * one stream per GPU, n-step chaining as n launches on that stream with ping-pong source
 * buffers (no host round trip, no per-step device copy), lazy AoS download.
 *
 * Multi-GPU: targets are sharded by tiles over all GPUs of all cooperating processes; after each
 * step the new positions of the massive particles are re-replicated -- by peer stores from the
 * kernel epilogue when the GPUs belong to one process, by an NCCL all-gather (a group of
 * ncclBroadcast with ragged counts) otherwise -- overlapped with the massless targets' kernel.
 */
#include <algorithm>
#include <cstring>
#include <vector>

#include "balanced_kernel.cuh"
#include "exact_kernel.cuh"
#include "layout_kernels.cuh"
#include "nbody_sim.h"
#include "nccl_dyn.h"
#include "persist_kernel.cuh"
#include "render_kernels.cuh"
#include "resident_kernel.cuh"
#include "small_kernel.cuh"
#include "split_kernel.cuh"
#include "step_kernel.cuh"
#include "util.cuh"

namespace nb {

/* ---- kernel variants ------------------------------------------------------------------ */

struct Variant {
    const char *name;
    int threads, T, tjp, ns, minb, unr;
    double rel_cost; /* measured time per pair relative to the best variant at large N (0: not a plan candidate) */
    bool cluster;    /* the nslab CTAs of a tile are one thread-block cluster (cluster_reduce.cuh): nslab <= 16 */
    size_t smem;
    const void *func;
    void (*launch)(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl);
    /* the same kernel walking a group of slabs per CTA (step_kernel GROUPED); null where it is not built */
    size_t smem_grouped;
    const void *func_grouped;
    void (*launch_grouped)(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl);
    /* n chained steps in one cooperative launch (persist_kernel.cuh); null where it is not built */
    const void *func_persist;
    void (*launch_persist)(uint32_t grid, cudaStream_t st, const PersistParams &pp);
    /* small-World kernel (small_kernel.cuh): a tile is 32 * T targets, its slabs are split over the warps of ONE
     * CTA, all sources sit in shared memory (sized per World at launch) */
    bool small;
    /* the reference's CPU arithmetic reproduced bit for bit (exact_kernel.cuh): one thread per target, no slabs */
    bool exact;
    /* mid-size Worlds (balanced_kernel.cuh): rows of 32 * T targets x slabs, shared out evenly over one CTA per SM */
    bool balanced;
};

template <int LANES>
static void launch_exact(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kExactThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    unsigned na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    NB_CUDA(cudaLaunchKernelEx(&cfg, exact_kernel<LANES>, p));
}

template <int WARPS, int T, int UNR, bool TL = false>
static void launch_small(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(WARPS * 32);
    cfg.dynamicSmemBytes = small_smem_bytes(p.src_packs_pad, p.nslab, T);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    unsigned na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    NB_CUDA(cudaLaunchKernelEx(&cfg, small_world_kernel<WARPS, T, UNR, TL>, p));
}

/* GRID CTAs (at most one per SM) share out p.fold_tiles rows x p.nslab slabs */
template <int CW, int T, int UNR, bool TL = false>
static void launch_balanced(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3((CW + 1) * 32); /* CW compute warps and the finisher */
    cfg.dynamicSmemBytes = balanced_smem_bytes(p.src_packs_pad, p.nslab, T, p.fold_tiles, grid);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    unsigned na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    NB_CUDA(cudaLaunchKernelEx(&cfg, balanced_kernel<CW, T, UNR, TL>, p));
}

/* GRID = tiles x CS CTAs in clusters of CS: the CTAs of a cluster share a tile's slabs (split_kernel.cuh) */
template <int WARPS, int T, int UNR>
static void launch_small_split(uint32_t grid, uint32_t cs, cudaStream_t st, const StepParams &p, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(WARPS * 32);
    cfg.dynamicSmemBytes = small_smem_bytes(p.src_packs_pad, p.nslab, T);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    unsigned na = 0;
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cs;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    na++;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    NB_CUDA(cudaLaunchKernelEx(&cfg, small_split_kernel<WARPS, T, UNR>, p));
}

template <int THREADS, int T, int TJP, int NS, int MINB, int UNR>
static void launch_persistent(uint32_t grid, cudaStream_t st, const PersistParams &pp) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = step_smem_bytes<TJP, NS>();
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative; /* all CTAs co-resident, or the launch fails: they meet in a grid barrier */
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    NB_CUDA(cudaLaunchKernelEx(&cfg, step_kernel_persistent<THREADS, T, TJP, NS, MINB, UNR>, pp));
}

template <int THREADS, int T, int TJP, int NS, int MINB, int UNR, int FORM, bool GROUPED = false>
static void launch_variant(uint32_t grid, cudaStream_t st, const StepParams &p, bool pdl) {
    /* Programmatic dependent launch: the next step's CTAs may become resident while this step's last
     * CTAs drain; they run their prologue and block in griddepcontrol.wait until this grid has
     * completed and flushed (step_kernel.cuh).  Hides launch latency between chained steps. */
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = GROUPED ? step_smem_bytes_grouped<THREADS, T, TJP, NS>() : step_smem_bytes<TJP, NS>();
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    unsigned na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    if (FORM == kFormCluster) { /* one cluster per target tile: its slab CTAs */
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = p.nslab;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    NB_CUDA(cudaLaunchKernelEx(&cfg, step_kernel<THREADS, T, TJP, NS, MINB, UNR, FORM, GROUPED>, p));
}

#define NB_VARIANT(NAME, THREADS, T, TJP, NS, MINB, UNR, FORM, REL)                                   \
    {                                                                                                 \
        NAME, THREADS, T, TJP, NS, MINB, UNR, REL, FORM == kFormCluster, step_smem_bytes<TJP, NS>(),  \
            (const void *)step_kernel<THREADS, T, TJP, NS, MINB, UNR, FORM>,                          \
            launch_variant<THREADS, T, TJP, NS, MINB, UNR, FORM>, 0, nullptr, nullptr, nullptr, nullptr, false, \
            false, false                                                                                     \
    }
/* a plan candidate: also built as a GROUPED kernel for Worlds whose per-slab partials exceed the budget */
#define NB_VARIANT_G(NAME, THREADS, T, TJP, NS, MINB, UNR, REL, PERSIST_FUNC, PERSIST_LAUNCH)                \
    {                                                                                                        \
        NAME, THREADS, T, TJP, NS, MINB, UNR, REL, false, step_smem_bytes<TJP, NS>(),                        \
            (const void *)step_kernel<THREADS, T, TJP, NS, MINB, UNR, kFormPacked>,                          \
            launch_variant<THREADS, T, TJP, NS, MINB, UNR, kFormPacked>,                                     \
            step_smem_bytes_grouped<THREADS, T, TJP, NS>(),                                                  \
            (const void *)step_kernel<THREADS, T, TJP, NS, MINB, UNR, kFormPacked, true>,                    \
            launch_variant<THREADS, T, TJP, NS, MINB, UNR, kFormPacked, true>, PERSIST_FUNC, PERSIST_LAUNCH, \
            false, false, false                                                                              \
    }
/* the small-World kernel: WARPS warps per CTA, T targets per lane; TJP only rounds the source padding */
#define NB_VARIANT_SMALL(NAME, WARPS, T, UNR) NB_VARIANT_SMALL_TL(NAME, WARPS, T, UNR, false)
#define NB_VARIANT_SMALL_TL(NAME, WARPS, T, UNR, TL)                                                         \
    {                                                                                                        \
        NAME, WARPS * 32, T, 64, 0, 1, UNR, 0.0, false, 0, (const void *)small_world_kernel<WARPS, T, UNR, TL>, \
            launch_small<WARPS, T, UNR, TL>, 0, nullptr, nullptr, nullptr, nullptr, true, false, false       \
    }
/* the balanced kernel of mid-size Worlds: CW compute warps + 1 finisher warp per CTA, T targets per lane, one CTA per SM */
#define NB_VARIANT_BALANCED(NAME, CW, T, UNR, TL)                                                            \
    {                                                                                                        \
        NAME, (CW + 1) * 32, T, 64, 0, 1, UNR, 0.0, false, 0, (const void *)balanced_kernel<CW, T, UNR, TL>, \
            launch_balanced<CW, T, UNR, TL>, 0, nullptr, nullptr, nullptr, nullptr, false, false, true       \
    }
/* the reference-exact verification kernel for the reference build with LANES SIMD lanes */
#define NB_VARIANT_EXACT(NAME, LANES)                                                                        \
    {                                                                                                        \
        NAME, kExactThreads, 1, 64, 0, 1, 1, 0.0, false, 0, (const void *)exact_kernel<LANES>,               \
            launch_exact<LANES>, 0, nullptr, nullptr, nullptr, nullptr, false, true, false                   \
    }
/* the small-World candidate is also built as the persistent multi-step kernel (persist_kernel.cuh) */
#define NB_PERSIST_FUNC(THREADS, T, TJP, NS, MINB, UNR) (const void *)step_kernel_persistent<THREADS, T, TJP, NS, MINB, UNR>
#define NB_PERSIST_LAUNCH(THREADS, T, TJP, NS, MINB, UNR) launch_persistent<THREADS, T, TJP, NS, MINB, UNR>

static const Variant kVariants[] = {
    /* plan candidates, with their measured cost per pair relative to t512x4 (profiles/r01_tune_sweep*.txt) */
    NB_VARIANT_G("t512x4", 512, 4, 128, 4, 1, 2, 1.000, nullptr, nullptr), /* 2048 targets per CTA, 1 CTA/SM  */
    NB_VARIANT_G("t256x4", 256, 4, 128, 4, 2, 2, 1.010, nullptr, nullptr), /* 1024 targets per CTA            */
    NB_VARIANT_G("t128x4", 128, 4, 128, 4, 4, 2, 1.025, nullptr, nullptr), /* 512 targets per CTA             */
    NB_VARIANT_G("t128x2", 128, 2, 64, 4, 6, 2, 1.065, NB_PERSIST_FUNC(128, 2, 64, 4, 6, 2),
                 NB_PERSIST_LAUNCH(128, 2, 64, 4, 6, 2)), /* 256 targets per CTA, small N */
    /* selectable with NBODY_VARIANT only */
    NB_VARIANT("t256x2", 256, 2, 128, 4, 3, 2, kFormPacked, 0.0),
    NB_VARIANT("t128x8", 128, 8, 128, 4, 2, 1, kFormPacked, 0.0),
    NB_VARIANT("t64x4", 64, 4, 64, 4, 8, 2, kFormPacked, 0.0),
    /* small Worlds: slabs split over the warps of one CTA, sources resident in shared memory (small_kernel.cuh);
     * chosen by small_variant_for, or with NBODY_VARIANT */
    NB_VARIANT_SMALL("w32x1", 32, 1, 2), /* planned: N <= 32 SMs         */
    NB_VARIANT_SMALL("w16x2", 16, 2, 2), /* planned: N <= 64 SMs         */
    NB_VARIANT_SMALL("w16x3", 16, 3, 2), /* planned: 0.78 x 96 SMs <= N <= 96 SMs   */
    NB_VARIANT_SMALL("w16x4", 16, 4, 2), /* planned: 0.78 x 128 SMs <= N <= 128 SMs */
    NB_VARIANT_SMALL("w16x1", 16, 1, 2), /* NBODY_VARIANT only (the sweep's other candidates) */
    NB_VARIANT_SMALL("w8x1", 8, 1, 2),
    NB_VARIANT_SMALL_TL("w16x4t", 16, 4, 2, true), /* with the per-warp timeline (tools/step_timeline.py) */
    NB_VARIANT_SMALL_TL("w16x3t", 16, 3, 2, true),
    NB_VARIANT_SMALL_TL("w32x1t", 32, 1, 2, true),
    /* mid-size Worlds: (row, slab) units shared out evenly over one CTA per SM (balanced_kernel.cuh) */
    NB_VARIANT_BALANCED("b16x1", 16, 1, 2, false),
    NB_VARIANT_BALANCED("b16x2", 16, 2, 2, false),
    NB_VARIANT_BALANCED("b16x3", 16, 3, 2, false),
    NB_VARIANT_BALANCED("b16x4", 16, 4, 2, false),
    NB_VARIANT_BALANCED("b16x1t", 16, 1, 2, true), /* with the per-warp timeline (tools/step_timeline.py) */
    NB_VARIANT_BALANCED("b16x2t", 16, 2, 2, true),
    NB_VARIANT_BALANCED("b16x4t", 16, 4, 2, true),
    /* verification mode: the reference's AVX / SSE / scalar CPU arithmetic bit for bit (exact_kernel.cuh);
     * NBODY_EXACT=avx|sse|scalar, NB_SIM_FLAG_EXACT (AVX), or NBODY_VARIANT */
    NB_VARIANT_EXACT("ref8", 8),
    NB_VARIANT_EXACT("ref4", 4),
    NB_VARIANT_EXACT("ref1", 1),
    /* slab reduction through distributed shared memory: bit-identical, measured slower (cluster_reduce.cuh) */
    NB_VARIANT("c128x2", 128, 2, 64, 4, 6, 2, kFormCluster, 0.0),
    NB_VARIANT("c128x4", 128, 4, 128, 4, 4, 2, kFormCluster, 0.0),
};
static const int kNumVariants = sizeof(kVariants) / sizeof(kVariants[0]);

/* ---- geometry ---------------------------------------------------------------------------- */

struct Plan {
    const Variant *var;
    uint32_t tile;           /* targets per tile                                    */
    uint32_t ntiles;         /* all target tiles                                    */
    uint32_t mtiles;         /* tiles that contain at least one massive particle    */
    uint32_t n_pad;          /* ntiles * tile                                       */
    uint32_t src_packs;      /* ceil(mass_len / 4)                                  */
    uint32_t src_packs_pad;  /* rounded up to the smem tile                         */
    uint32_t slab_packs, nslab; /* the slabs: they fix the summation order                              */
    uint32_t group_slabs;       /* consecutive slabs one CTA walks (1: one CTA per slab)                */
    uint32_t ngroups;           /* CTAs per target tile = ceil(nslab / group_slabs) = partials per target */
    uint32_t csplit;            /* small Worlds with few tiles: CTAs (one cluster) that share a tile's slabs    */
                                /* (split_kernel.cuh); 1: one CTA per tile.  Not part of the summation order   */
    int ctas_per_sm;
};

static uint32_t ceil_div(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

/*
 * Work plan = (kernel variant, slab length), chosen together by a timing model fitted to sweeps on
 * B200 (tools/tune.cu sweep; fit within ~3 % for N >= 50K, ~10 % below):
 *
 *   a grid of C = tiles * nslab CTAs runs as full waves of `slots` = SMs * CTAs/SM plus one partial
 *   wave; a CTA computes tile * slab pairs; an SM delivers `kPairsPerClk` pairs per clock when 16 or
 *   more warps are resident and a fraction 0.7 + 0.3 * warps/16 of that below; every wave pays a fixed
 *   prologue/epilogue cost and every launch a fixed latency.
 *
 * The plan fixes the summation order (tile size, slab boundaries), so it must not depend on how many
 * GPUs run the World: the model is evaluated for the World split over 1, 2, 4 and 8 GPUs and the
 * weighted total GPU time is minimised (1 and 8 GPUs weigh most).  Results are bit-identical across
 * GPU counts because of this.
 */
static const double kPairsPerClk = 12.0;    /* per SM at full occupancy, best variant (65.5 % of FP32 peak) */
static const double kWaveFixedClk = 8000.0; /* per wave of CTAs: first TMA round trip, partial store, ticket, reduce */
static const double kLaunchClk = 12000.0;   /* per launch */
static const uint32_t kMinTargetsPerGpu = 32768; /* below this a GPU is latency-bound: not a split to plan for */
static const uint32_t kMinSlabs = 16;       /* accuracy floor, see choose_slabs */
static const uint32_t kShortWorldSources = 2048; /* ... except below this many sources, where the floor is */
static const uint32_t kMinSlabSources = 256;     /*     slabs of this length (chains of 128 terms)             */
static const double kReduceRoundClk = 1400.0; /* per 8 slabs: one dependent L2 round trip of the last-CTA reduction */
static const uint32_t kMaxSlabSources = 16384;
static const uint32_t kMinTiles = 48;       /* a variant needs this many target tiles to be a candidate */
static const uint32_t kMaxClusterSlabs = 16; /* largest (non-portable) thread-block cluster on sm_100 */

static double model_clocks(const Variant &v, uint32_t tiles, uint32_t nslab, uint32_t slab_sources, int sm_count,
                           bool short_world) {
    const double ctas = (double)tiles * nslab;
    const double slots = (double)v.minb * sm_count;
    const int warps = v.threads / 32;
    const double cta_pairs = (double)v.threads * v.T * slab_sources * v.rel_cost;
    auto occ_eff = [](int w) { return w >= 16 ? 1.0 : 0.7 + 0.3 * w / 16.0; };
    const double full = std::floor(ctas / slots), rem = ctas - full * slots;
    double t = kLaunchClk;
    /* the tail of the launch is a last-CTA reduction: ceil(nslab / 8) dependent rounds of partial loads (measured at
     * N = 4 000 with 16 tiles: 8 slabs 8.0 us, 13-16 slabs 8.8 us, profiles/r01_sweep_plan_smallN.txt; applied only
     * to the short Worlds it was measured on -- larger ones keep the plans their parity runs were made with) */
    if (nslab > 1 && short_world) t += kReduceRoundClk * std::ceil(nslab / 8.0);
    if (full > 0) t += full * (cta_pairs * v.minb / (kPairsPerClk * occ_eff(v.minb * warps)) + kWaveFixedClk);
    if (rem > 0) {
        const int k = (int)std::ceil(rem / sm_count);
        t += cta_pairs * k / (kPairsPerClk * occ_eff(k * warps)) + kWaveFixedClk;
    }
    return t;
}

static void set_geometry(Plan &pl, const Variant *v, uint32_t N, uint32_t M) {
    pl.var = v;
    pl.ctas_per_sm = v->minb;
    pl.tile = (v->small || v->balanced) ? 32u * (uint32_t)v->T : (uint32_t)v->threads * v->T;
    pl.ntiles = ceil_div(N, pl.tile);
    pl.mtiles = ceil_div(M, pl.tile);
    pl.n_pad = pl.ntiles * pl.tile;
    pl.src_packs = ceil_div(M, 4);
    pl.src_packs_pad = std::max(1u, ceil_div(pl.src_packs, (uint32_t)v->tjp)) * v->tjp;
}

/* Model time of a plan with NS CTAs per target tile of SP packs each, weighted over the GPU counts a World may be
 * split over (GPU-clocks summed over GPUs). */
static double weighted_cost(const Plan &pl, uint32_t N, int sm_count, uint32_t ns, uint32_t sp) {
    static const int kGpus[4] = {1, 2, 4, 8};
    static const double kWeight[4] = {4.0, 1.0, 1.0, 2.0};
    double cost = 0.0, wsum = 0.0;
    for (int q = 0; q < 4; q++) {
        const uint32_t g = (uint32_t)kGpus[q];
        /* a split is only worth planning for when every GPU keeps a throughput-bound share */
        if (g > 1 && (pl.ntiles < 2 * g || N / g < kMinTargetsPerGpu)) break;
        cost += kWeight[q] * g * model_clocks(*pl.var, ceil_div(pl.ntiles, g), ns, sp * 4, sm_count,
                                              pl.src_packs * 4 < kShortWorldSources);
        wsum += kWeight[q];
    }
    return cost / wsum;
}

/* Bytes the slab partials of ONE GPU running the whole World may take (nslab x targets x 8 B).  A constant, not
 * the free memory of the device at hand: the plan fixes the summation order and must be a function of (N, M)
 * only.  NBODY_PART_BUDGET (bytes) overrides it -- for tests, which force slab groups at small N with it. */
static const uint64_t kPartBudgetBytes = 8ull << 30; /* N <= 4M (4.3 GB) keeps one CTA per slab, as measured */
static uint64_t part_budget_bytes() {
    const char *e = getenv("NBODY_PART_BUDGET");
    if (e && *e) return std::max<uint64_t>(1, strtoull(e, nullptr, 10));
    return kPartBudgetBytes;
}

/*
 * Slab groups: bound the partial buffer.  The slab count grows with M (no accumulator chain longer than
 * kMaxSlabSources / 2 terms), so one partial per (slab, target) is nslab x N x 8 B = O(N x M): 310 MB at N = 1M,
 * 4.3 GB at 4M, 69 GB at 16M.  Past the budget (the GROUPED build's pair loop is 1.7 % slower than the ungrouped
 * one even with its own best source order -- profiles/r02_schedule_lottery_grouped.txt -- so it is not used where
 * the plain buffer is affordable) a CTA walks `group_slabs` consecutive slabs and adds their sums in
 * slab order itself (step_kernel GROUPED); only ngroups = ceil(nslab / group_slabs) partials per target reach
 * memory.  Slabs become whole source tiles for this; the number of groups is chosen by the same timing model as
 * the slab count.  Returns the model cost of the final plan.
 */
static double group_slabs(Plan &pl, uint32_t N, int sm_count, double cost) {
    pl.group_slabs = 1;
    pl.ngroups = pl.nslab;
    const uint64_t per_partial = (uint64_t)pl.n_pad * sizeof(float2);
    const uint64_t budget = part_budget_bytes();
    if (pl.nslab <= 1 || pl.var->func_grouped == nullptr || (uint64_t)pl.nslab * per_partial <= budget) return cost;
    const uint32_t tjp = (uint32_t)pl.var->tjp;
    const uint32_t sp = std::max(tjp, pl.slab_packs / tjp * tjp);
    const uint32_t ns = ceil_div(pl.src_packs, sp);
    const uint32_t max_groups = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(ns, budget / per_partial));
    double best = 1e300;
    uint32_t best_gs = ns;
    for (uint32_t g = max_groups; g >= 1; g--) { /* on near ties: more, shorter CTAs */
        const uint32_t gs = ceil_div(ns, g);
        const uint32_t groups = ceil_div(ns, gs);
        if (groups != g) continue;
        const double c = weighted_cost(pl, N, sm_count, groups, gs * sp);
        if (c < best * 0.999) {
            best = c;
            best_gs = gs;
        }
    }
    pl.slab_packs = sp;
    pl.nslab = ns;
    pl.group_slabs = best_gs;
    pl.ngroups = ceil_div(ns, best_gs);
    return best;
}

/* Best slab count for one variant; returns the weighted model time (GPU-clocks summed over GPUs). */
static double choose_slabs_ungrouped(Plan &pl, uint32_t N, int sm_count, uint32_t forced_slab_len) {
    const uint32_t unr = 4; /* slab granularity in packs: a multiple of every variant's unroll */
    auto weighted = [&](uint32_t ns, uint32_t sp) { return weighted_cost(pl, N, sm_count, ns, sp); };
    if (pl.src_packs == 0) {
        pl.slab_packs = unr;
        pl.nslab = 1;
        return weighted(1, unr);
    }
    const uint32_t cap = pl.var->cluster ? kMaxClusterSlabs : 65536u; /* slabs per tile (grouped beyond the budget) */
    if (forced_slab_len) {
        pl.slab_packs = std::max(unr, ceil_div(ceil_div(forced_slab_len, 4), unr) * unr);
        pl.slab_packs = std::max(pl.slab_packs, ceil_div(ceil_div(pl.src_packs, cap), unr) * unr);
        pl.nslab = ceil_div(pl.src_packs, pl.slab_packs);
        return weighted(pl.nslab, pl.slab_packs);
    }
    const uint32_t min_packs = 32; /* >= 128 sources per slab */
    const uint32_t max_nslab = std::max(1u, std::min(cap, pl.src_packs / min_packs));
    /* Accuracy floor: a register accumulator chain sums slab_len/2 terms, the reference's AVX build
     * M/8 (ref sim_cpu.c:165-189).  At least kMinSlabs slabs keeps our chains 4x shorter than the
     * reference's, so the CUDA result stays closer to FP64 than the reference is whatever the model
     * would prefer (measured at N = 50K: 6 slabs p99 1.5e-5 vs FP64, 45 slabs 2.2e-6, reference 1.9e-5);
     * and no chain is longer than kMaxSlabSources/2 terms however large the World is.  Below
     * kShortWorldSources the floor relaxes to slabs of kMinSlabSources: chains of 128 terms are short whatever
     * their number (N = 4 003 fixture, FP32 emulation of both plans: p99 against FP64 5.8e-7 -> 8.0e-7). */
    const uint32_t floor_slabs = pl.src_packs * 4 < kShortWorldSources
                                     ? std::max(1u, ceil_div(pl.src_packs * 4, kMinSlabSources))
                                     : kMinSlabs;
    const uint32_t min_nslab = std::min(std::max(floor_slabs, ceil_div(pl.src_packs * 4, kMaxSlabSources)), max_nslab);
    double best = 1e300;
    uint32_t best_n = min_nslab;
    for (uint32_t n = min_nslab; n <= max_nslab; n++) {
        const uint32_t sp = ceil_div(ceil_div(pl.src_packs, n), unr) * unr;
        if (ceil_div(pl.src_packs, sp) != n) continue;
        const double cost = weighted(n, sp);
        if (cost < best * 0.999) {
            best = cost;
            best_n = n;
        }
    }
    pl.slab_packs = ceil_div(ceil_div(pl.src_packs, best_n), unr) * unr;
    pl.nslab = ceil_div(pl.src_packs, pl.slab_packs);
    if (best >= 1e300) best = weighted(pl.nslab, pl.slab_packs); /* no count in range divides evenly */
    return best;
}

/* Slabs of a small World (small_kernel.cuh): the warps of the tile's one CTA share them out, so equal lengths,
 * one per warp -- or k per warp once a slab would exceed 512 sources (chains of 256 terms).  The slab sums never
 * leave shared memory: no partial buffer, no groups. */
static void choose_slabs_small(Plan &pl, uint32_t forced_slab_len) {
    const uint32_t unr = 4;
    const uint32_t warps = (uint32_t)pl.var->threads / 32;
    pl.group_slabs = 1;
    pl.ngroups = 1;
    if (pl.src_packs == 0) {
        pl.slab_packs = unr;
        pl.nslab = 1;
        return;
    }
    if (forced_slab_len) {
        pl.slab_packs = std::max(unr, ceil_div(ceil_div(forced_slab_len, 4), unr) * unr);
    } else {
        const uint32_t per_warp = ceil_div(pl.src_packs, warps * 128u);
        pl.slab_packs = std::max(unr, ceil_div(ceil_div(pl.src_packs, warps * per_warp), unr) * unr);
    }
    pl.nslab = ceil_div(pl.src_packs, pl.slab_packs);
}

/* Slabs of a mid-size World (balanced_kernel.cuh): a CTA gets at most ceil(rows * nslab / CTAs) units and cuts them
 * into WARPS contiguous runs, so a warp walks ceil(that / WARPS) units at most: the step takes as long as the longest
 * run.  A unit boundary inside a run costs next to nothing (a slab-sum store), so slabs are SHORT -- ten or more units
 * per warp, runs that differ by one unit in ten -- down to 32 sources, as far as the unit sums fit in shared memory
 * beside the sources.  Shorter chains only help the accuracy.  Cost of a plan = longest run x (slab + fixed). */
static const double kBalancedUnitFixedPacks = 0.5; /* per unit, in packs of work: slab-sum store, chain restart */
static const size_t kBalancedMaxSmem = 220u << 10;
static void choose_slabs_balanced(Plan &pl, int sm_count, uint32_t forced_slab_len) {
    const uint32_t unr = 4;
    const uint32_t warps = (uint32_t)pl.var->threads / 32 - 1; /* compute warps: one warp is the finisher */
    pl.group_slabs = 1;
    if (pl.src_packs == 0) {
        pl.slab_packs = unr;
        pl.nslab = pl.ngroups = 1;
        return;
    }
    if (forced_slab_len) {
        pl.slab_packs = std::max(unr, ceil_div(ceil_div(forced_slab_len, 4), unr) * unr);
    } else {
        /* the shortest slabs within 1 % of the cheapest plan */
        double best = 1e300;
        uint32_t best_sp = 128;
        const uint32_t n_lo = std::max(1u, ceil_div(pl.src_packs, 128u)), n_hi = std::max(n_lo, pl.src_packs / 8u);
        for (int pass = 0; pass < 2; pass++) {
            for (uint32_t n = n_lo; n <= n_hi; n++) {
                const uint32_t sp = ceil_div(ceil_div(pl.src_packs, n), unr) * unr;
                if (ceil_div(pl.src_packs, sp) != n) continue;
                const uint32_t units = pl.ntiles * n, ctas = std::min<uint32_t>((uint32_t)sm_count, units);
                if (n > n_lo && balanced_smem_bytes(pl.src_packs_pad, n, pl.var->T, pl.ntiles, ctas) > kBalancedMaxSmem) break;
                const uint32_t run = ceil_div(balanced_geom(pl.ntiles, n, ctas).upc_max, warps);
                const double cost = run * ((double)sp + kBalancedUnitFixedPacks);
                if (pass == 0) best = std::min(best, cost);
                else if (cost <= best * 1.01) best_sp = sp;
            }
        }
        pl.slab_packs = best_sp;
    }
    pl.nslab = ceil_div(pl.src_packs, pl.slab_packs);
    pl.ngroups = pl.nslab; /* one partial per (slab, target) in memory, for the rows that straddle two CTAs */
}

static size_t balanced_plan_smem(const Plan &pl, int sm_count) {
    const uint32_t units = pl.ntiles * pl.nslab;
    return balanced_smem_bytes(pl.src_packs_pad, pl.nslab, pl.var->T, pl.ntiles, std::min<uint32_t>((uint32_t)sm_count, units));
}

static double choose_slabs(Plan &pl, uint32_t N, int sm_count, uint32_t forced_slab_len) {
    if (pl.var->exact) { /* the reference's order: LANES chains over all sources, no slabs */
        pl.slab_packs = std::max(4u, pl.src_packs_pad);
        pl.nslab = 1;
        pl.group_slabs = 1;
        pl.ngroups = 1;
        return 0.0;
    }
    if (pl.var->small) {
        choose_slabs_small(pl, forced_slab_len);
        return 0.0;
    }
    if (pl.var->balanced) {
        choose_slabs_balanced(pl, sm_count, forced_slab_len);
        return 0.0;
    }
    const double cost = choose_slabs_ungrouped(pl, N, sm_count, forced_slab_len);
    return group_slabs(pl, N, sm_count, cost);
}

static const Variant *find_variant(const char *name) {
    for (int i = 0; i < kNumVariants; i++)
        if (strcmp(name, kVariants[i].name) == 0) return &kVariants[i];
    return nullptr;
}

/*
 * Which Worlds take the small-World kernel (small_kernel.cuh), and with how many targets per lane.  One CTA per
 * tile of 32 * T targets and no more CTAs than SMs (a second CTA on an SM doubles that SM's work: measured cliff),
 * so T = ceil(N / (32 * SMs)), up to 4.  Measured against the (tile, slab)-per-CTA kernel on B200
 * (profiles/r02_small_world_variants.txt, us per step): N = 1 000 3.6 vs 6.9; 4 000 5.7 vs 8.1; 4 736 6.4 vs 10.6;
 * 8 000 14.3 vs 18.4; 9 472 16.4 vs 23.2 (T <= 2: always faster); T = 3 and 4 win only when most SMs get a CTA
 * (12 000 28.6 vs 30.9, 14 208 32.8 vs 39.2, 18 944 56.2 vs 63.3, but 10 000 24.3 vs 22.5 and 15 000 a tie), hence
 * the utilisation floor.  Best warp count per T from the same sweep: 32 warps at T = 1, 16 above.
 * NBODY_SMALL_MAX_N caps N (0: never use the small-World kernel).
 */
static const size_t kSmallWorldMaxSmem = 200u << 10;
static const double kSmallWorldMinUtil = 0.78; /* T >= 3: CTAs / SMs */
static const Variant *small_variant_for(uint32_t N, int sm_count) {
    const char *e = getenv("NBODY_SMALL_MAX_N");
    if (e && *e && N > (uint32_t)strtoul(e, nullptr, 10)) return nullptr;
    const uint32_t T = ceil_div(N, 32u * (uint32_t)sm_count);
    if (T == 0 || T > 4) return nullptr;
    if (T >= 3 && (double)ceil_div(N, 32u * T) < kSmallWorldMinUtil * sm_count) return nullptr;
    const char *name = T == 1 ? "w32x1" : T == 2 ? "w16x2" : T == 3 ? "w16x3" : "w16x4";
    return find_variant(name);
}

/*
 * Which Worlds take the balanced kernel (balanced_kernel.cuh): the mid sizes the small-World kernel leaves out -- T = 3, 4
 * with too few tiles for the SMs, and T = 5, 6 -- where the (tile, slab)-per-CTA kernel runs a fraction of a wave too
 * many.  Measured on B200 (profiles/r02_balanced_v4_bands.txt, us per step, (tile, slab) kernel -> b16x2): N = 9 500
 * 21.5 -> 19.5; 10 500 25.0 -> 22.5; 11 000 27.4 -> 23.8; 14 700 42.2 -> 39.5; 19 000 65.4 -> 62.3; 20 000 72.0 -> 68.1;
 * 22 000 84.1 -> 81.8; 24 000 98.0 -> 95.3; a tie at 10 000, 14 300, 21 000 and 23 000; a loss from 25 000 up.
 * NBODY_SMALL_MAX_N caps it like the small-World kernel; NBODY_NO_BALANCED=1 switches it off.
 */
static const uint32_t kBalancedMaxTargets = 24000;
static const Variant *balanced_variant_for(uint32_t N, int sm_count) {
    const char *e = getenv("NBODY_SMALL_MAX_N");
    if (e && *e && N > (uint32_t)strtoul(e, nullptr, 10)) return nullptr;
    if (getenv("NBODY_NO_BALANCED") && atoi(getenv("NBODY_NO_BALANCED"))) return nullptr;
    if (N <= 64u * (uint32_t)sm_count || N > kBalancedMaxTargets) return nullptr; /* T <= 2: the small-World kernel */
    return find_variant("b16x2");
}

static const Variant *forced_variant() {
    if (const char *x = getenv("NBODY_EXACT")) { /* reference-exact verification mode, by SIMD build of the reference */
        if (!strcmp(x, "avx") || !strcmp(x, "8") || !strcmp(x, "1") || !strcmp(x, "on")) return find_variant("ref8");
        if (!strcmp(x, "sse") || !strcmp(x, "4")) return find_variant("ref4");
        if (!strcmp(x, "scalar")) return find_variant("ref1");
        if (*x && strcmp(x, "0") && strcmp(x, "off")) NB_FAIL("NBODY_EXACT=%s: expected avx, sse or scalar", x);
    }
    const char *e = getenv("NBODY_VARIANT");
    if (!e || !*e) return nullptr;
    if (const Variant *v = find_variant(e)) return v;
    NB_FAIL("NBODY_VARIANT=%s is not a known kernel variant", e);
}

/* ---- pipeline state ------------------------------------------------------------------ */

enum ExchangeMode { kExchangeNone = 0, kExchangePeerStore = 1, kExchangeNccl = 2 };

struct Shard { /* one GPU driven by this process */
    int device = 0;
    int global_index = 0; /* rank * local_gpus + local index */
    cudaStream_t stream = nullptr, comm_stream = nullptr;
    bool own_stream = true;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_massive = nullptr, ev_comm = nullptr, ev_step = nullptr;
    Particle *aos = nullptr;
    float4 *src_xy[2] = {nullptr, nullptr};
    float4 *src_gm = nullptr;
    float2 *pos = nullptr, *vel = nullptr, *acc = nullptr, *part = nullptr;
    float *radius = nullptr;
    unsigned *tile_cnt = nullptr;
    unsigned *grid_bar = nullptr; /* arrival counter of the persistent kernel's grid barrier */
    unsigned bar_base = 0;        /* its value once everything enqueued so far has run          */
    double *scratch64 = nullptr;
    uint32_t a0 = 0, a1 = 0; /* massive target tiles [a0, a1)  */
    uint32_t b0 = 0, b1 = 0; /* massless target tiles [b0, b1) */
    NcclComm comm = nullptr;
};

}  // namespace nb

namespace nb {
static const uint32_t kGraphSteps = 20;        /* steps per CUDA-graph chunk (even)                          */
static const uint32_t kGraphMaxTargets = 4736; /* Worlds up to this size replay chained steps as graph chunks */
static const uint32_t kResidentMaxTargets = 640; /* Worlds up to this size run chained steps in one resident cluster: */
                                                 /* beyond ~750 the 148-SM paths win (resident_kernel.cuh)           */
}  // namespace nb

/* A chunk of kGraphSteps chained steps of a small World as one instantiated CUDA graph (step_async): valid for one
 * dt and for the source buffer the chunk starts from (the launches' arguments are baked into the graph). */
struct StepGraph {
    cudaGraphExec_t exec = nullptr;
    float dt = 0.f;
};

struct SimPipeline {
    WorldData data;
    nb::Plan plan;
    std::vector<nb::Shard> shards;
    int rank = 0, world = 1, local_gpus = 1, total_shards = 1;
    int exchange = nb::kExchangeNone;
    bool overlap = true;
    bool pdl = true;          /* programmatic dependent launch between chained steps */
    bool allow_persist = false; /* n chained steps of a small World as one cooperative launch: opt-in, measured
                                   slower than n launches with programmatic dependent launch (persist_kernel.cuh) */
    uint32_t persist_ctas = 0; /* largest grid the persistent multi-step kernel may be launched with (0: never) */
    bool use_graph = false;    /* chained steps of this World are host-launch-bound: run them as CUDA-graph chunks */
    StepGraph graphs[2];       /* one per source buffer a chunk may start from */
    int cur = 0;              /* which src_xy buffer holds the current positions */
    bool has_data = false;
    bool inert = false;       /* total_len == 0: no device state, every call is a no-op (the reference's CPU
                                 path accepts an empty World, ref sim_cpu.c:109-111, and its world.c creates the
                                 pipeline eagerly, ref world.c:52) */
    bool timed = false;       /* ev_begin/ev_end recorded at least once          */
    uint64_t kernels_launched = 0;
    uint64_t persistent_launches = 0; /* of those, cooperative launches of the persistent multi-step kernel */
    uint64_t graph_launches = 0;      /* cudaGraphLaunch calls (each replays kGraphSteps step kernels) */
    uint32_t resident_ctas = 0;       /* cluster size of the resident multi-step kernel for this World (0: not used) */
    uint32_t resident_rows = 0;       /* target rows per CTA of that cluster */
    uint32_t resident_warps = 32;     /* warps per CTA: 8, 16 or 32, enough for the (row, slab) units of a CTA */
    uint64_t resident_launches = 0;
    cudaDeviceProp prop;
    /* tile ranges of every global shard, for the exchange: massive [A0,A1) */
    std::vector<uint32_t> all_a0, all_a1, all_b0, all_b1;
};

namespace nb {

static int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    if (!e || !*e) return dflt;
    return atoi(e);
}

/* The public entry points switch devices; the caller (a host program that may be on another device, e.g. torch)
 * gets its current device back when they return. */
struct DeviceGuard {
    int prev = -1;
    DeviceGuard() {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            (void)cudaGetLastError();
            prev = -1;
        }
    }
    ~DeviceGuard() {
        if (prev >= 0) (void)cudaSetDevice(prev);
    }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;
};

static void split_range(uint32_t first, uint32_t count, int parts, int idx, uint32_t *lo, uint32_t *hi) {
    *lo = first + (uint32_t)(((uint64_t)count * idx) / parts);
    *hi = first + (uint32_t)(((uint64_t)count * (idx + 1)) / parts);
}

template <typename Tp>
static Tp *dev_alloc(size_t count) {
    Tp *p = nullptr;
    NB_CUDA(cudaMalloc((void **)&p, std::max<size_t>(count, 1) * sizeof(Tp)));
    return p;
}

/* Geometry from sizes alone -- no device is touched, so the same plan can be computed (and
 * tested) on a machine without a GPU.  Occupancy is the variant's __launch_bounds__ promise. */
static void plan_order(uint32_t N, uint32_t M, int sm_count, uint32_t forced_slab, Plan &pl, const Variant *force) {
    if (const Variant *v = force ? force : forced_variant()) {
        set_geometry(pl, v, N, M);
        choose_slabs(pl, N, sm_count, forced_slab);
        if (v->small)
            NB_ASSERT(small_smem_bytes(pl.src_packs_pad, pl.nslab, v->T) <= kSmallWorldMaxSmem,
                      "NBODY_VARIANT=%s keeps all %u sources in shared memory: too many", v->name, M);
        if (v->balanced)
            NB_ASSERT(balanced_plan_smem(pl, sm_count) <= kBalancedMaxSmem,
                      "NBODY_VARIANT=%s keeps all %u sources in shared memory: too many", v->name, M);
        return;
    }
    if (const Variant *v = small_variant_for(N, sm_count)) {
        Plan cand{};
        set_geometry(cand, v, N, M);
        choose_slabs(cand, N, sm_count, forced_slab);
        if (small_smem_bytes(cand.src_packs_pad, cand.nslab, v->T) <= kSmallWorldMaxSmem) {
            pl = cand;
            return;
        }
    } else if (const Variant *b = balanced_variant_for(N, sm_count)) {
        Plan cand{};
        set_geometry(cand, b, N, M);
        choose_slabs(cand, N, sm_count, forced_slab);
        if (balanced_plan_smem(cand, sm_count) <= kBalancedMaxSmem) {
            pl = cand;
            return;
        }
    }
    double best = 1e300;
    Plan best_plan{};
    const Variant *smallest = nullptr;
    for (int i = 0; i < kNumVariants; i++)
        if (kVariants[i].rel_cost > 0.0) smallest = &kVariants[i]; /* candidates are listed big to small */
    for (int i = 0; i < kNumVariants; i++) {
        if (kVariants[i].rel_cost <= 0.0) continue;
        Plan cand{};
        set_geometry(cand, &kVariants[i], N, M);
        /* The model is validated where there are many tiles; with few, big tiles the per-tile costs it
         * does not see (target load, last-CTA reduction of nslab partials by `tiles` CTAs only) dominate
         * -- measured: N = 16K with 8 tiles of 2048 is 25 % slower than with 64 tiles of 256. */
        if (cand.ntiles < kMinTiles && &kVariants[i] != smallest) continue;
        const double cost = choose_slabs(cand, N, sm_count, forced_slab);
        if (cost < best * 0.995) { /* prefer the earlier (larger-tile) variant on near ties */
            best = cost;
            best_plan = cand;
        }
    }
    pl = best_plan;
}

/* The plan: the summation order (plan_order) and, where it does not touch that order, how it is launched.  A T = 1
 * small-World plan whose tiles would leave more than half of the SMs idle shares each tile's slabs over a cluster of
 * 2 ... 4 CTAs (split_kernel.cuh: the planned T = 1 kernel w32x1 only); NBODY_NO_SPLIT=1 keeps one CTA per tile. */
static const uint32_t kMaxClusterSplit = 4;
static const uint32_t kMinSplitPacks = 112; /* below ~450 sources the two cluster barriers cost what the shorter pair
                                               loop saves (profiles/r02_cluster_split.txt: N = 700 2.8 -> 3.0 us,
                                               N = 800 2.8 -> 2.8, N = 1 000 3.1 -> 3.0, N = 2 000 3.9 -> 3.5) */
static void plan_host(uint32_t N, uint32_t M, int sm_count, uint32_t forced_slab, Plan &pl,
                      const Variant *force = nullptr) {
    plan_order(N, M, sm_count, forced_slab, pl, force);
    pl.csplit = 1;
    if (!strcmp(pl.var->name, "w32x1") && pl.src_packs >= kMinSplitPacks && !env_int("NBODY_NO_SPLIT", 0)) {
        const uint32_t cs = std::min(std::min(kMaxClusterSplit, (uint32_t)sm_count / std::max(1u, pl.ntiles)), pl.nslab);
        if (cs >= 2) pl.csplit = cs;
    }
}

/* Which tiles shard q of S computes: an equal share of the tiles holding massive particles
 * (they feed the exchange) and an equal share of the purely massless tiles. */
static void shard_tiles(const Plan &pl, int S, int q, uint32_t *a0, uint32_t *a1, uint32_t *b0, uint32_t *b1) {
    split_range(0, pl.mtiles, S, q, a0, a1);
    split_range(pl.mtiles, pl.ntiles - pl.mtiles, S, q, b0, b1);
}

/* the kernel a plan runs: the variant itself, its GROUPED build when a CTA walks several slabs, or the cluster-split
 * build of the T = 1 small-World kernel */
static const void *plan_func(const Plan &pl) {
    if (pl.csplit > 1) return (const void *)small_split_kernel<32, 1, 2>;
    return pl.group_slabs > 1 ? pl.var->func_grouped : pl.var->func;
}
/* what the function is ALLOWED to be launched with: a per-function attribute shared by every pipeline of the process,
 * so for the small-World kernel (shared memory sized per World) it is the cap, not this World's need */
static size_t plan_smem_limit(const Plan &pl) {
    if (pl.var->small) return kSmallWorldMaxSmem;
    if (pl.var->balanced) return kBalancedMaxSmem;
    return pl.group_slabs > 1 ? pl.var->smem_grouped : pl.var->smem;
}
static size_t plan_smem(const Plan &pl, int sm_count) {
    if (pl.var->small) return small_smem_bytes(pl.src_packs_pad, pl.nslab, pl.var->T);
    if (pl.var->balanced) return balanced_plan_smem(pl, sm_count);
    return pl.group_slabs > 1 ? pl.var->smem_grouped : pl.var->smem;
}

static void make_plan(SimPipeline *sim, uint32_t forced_slab, bool exact) {
    Plan &pl = sim->plan;
    plan_host(sim->data.total_len, sim->data.mass_len, sim->prop.multiProcessorCount, forced_slab, pl,
              exact ? find_variant("ref8") : nullptr);
    NB_CUDA(cudaFuncSetAttribute(plan_func(pl), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan_smem_limit(pl)));
    if (pl.var->cluster) {
        NB_ASSERT(pl.nslab <= kMaxClusterSlabs, "cluster variant %s with %u slabs", pl.var->name, pl.nslab);
        if (pl.nslab > 8) NB_CUDA(cudaFuncSetAttribute(pl.var->func, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
    int occ = 0;
    NB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, plan_func(pl), pl.var->threads,
                                                          plan_smem(pl, sim->prop.multiProcessorCount)));
    NB_ASSERT(occ >= pl.var->minb, "kernel variant %s%s: %d CTAs/SM resident, %d planned", pl.var->name,
              pl.group_slabs > 1 ? " (grouped)" : "", occ, pl.var->minb);
    /* opt-in: n chained steps of a small World in one cooperative launch, only when the whole grid is co-resident */
    sim->persist_ctas = 0;
    if (pl.var->func_persist && pl.group_slabs == 1 && sim->total_shards == 1 && sim->allow_persist) {
        int coop = 0, pocc = 0, dev = 0;
        NB_CUDA(cudaGetDevice(&dev)); /* the pipeline's first device is current here */
        NB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        NB_CUDA(cudaFuncSetAttribute(pl.var->func_persist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.var->smem));
        NB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pocc, pl.var->func_persist, pl.var->threads, pl.var->smem));
        if (coop) sim->persist_ctas = (uint32_t)pocc * (uint32_t)sim->prop.multiProcessorCount;
    }
}

}  // namespace nb

/* ---- creation / destruction ---------------------------------------------------------- */

namespace nb {

static void particle_range_of_tiles(const SimPipeline *sim, uint32_t t0, uint32_t t1, uint32_t limit, uint32_t *lo,
                                    uint32_t *hi) {
    const uint64_t tile = sim->plan.tile;
    *lo = (uint32_t)std::min<uint64_t>(t0 * tile, limit);
    *hi = (uint32_t)std::min<uint64_t>(t1 * tile, limit);
}

/* One cluster of up to 16 CTAs runs all n steps of a call (resident_kernel.cuh): tiny Worlds with the T = 1 small-World
 * plan, on one GPU.  NBODY_RESIDENT_MAX_N moves the size limit (0: never). */
static void resident_launch_config(SimPipeline *sim, cudaLaunchConfig_t *cfg, cudaLaunchAttribute *attr, bool pdl) {
    memset(cfg, 0, sizeof *cfg);
    cfg->gridDim = dim3(sim->resident_ctas);
    cfg->blockDim = dim3(sim->resident_warps * 32);
    cfg->dynamicSmemBytes = resident_smem_bytes(sim->plan.src_packs_pad, sim->plan.nslab, sim->resident_rows);
    cfg->stream = sim->shards[0].stream;
    unsigned na = 0;
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = sim->resident_ctas;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    na++;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg->attrs = attr;
    cfg->numAttrs = na;
}

static const int kResidentUnroll = 4; /* packs per pair-loop iteration: the plan's slabs are multiples of 4 packs */
template <bool TL>
static const void *resident_func_tl(uint32_t warps) {
    return warps == 8    ? (const void *)resident_kernel<kResidentUnroll, 8, TL>
           : warps == 16 ? (const void *)resident_kernel<kResidentUnroll, 16, TL>
                         : (const void *)resident_kernel<kResidentUnroll, 32, TL>;
}
/* NBODY_RESIDENT_TL=1: the builds that record where the clocks of a step go (tools/resident_breakdown.py) */
static const void *resident_func(uint32_t warps) {
    return env_int("NBODY_RESIDENT_TL", 0) ? resident_func_tl<true>(warps) : resident_func_tl<false>(warps);
}

/* Host-only part of the decision (NbSimPlan reports it too): the cluster that would run chained steps of a World of
 * N particles under plan PL on one GPU -- CTAs (0: not this World), target rows per CTA, warps per CTA. */
static void resident_geometry(const Plan &pl, uint32_t N, uint32_t *ctas, uint32_t *rows_per_cta, uint32_t *warps) {
    const uint32_t max_n = (uint32_t)env_int("NBODY_RESIDENT_MAX_N", (int)kResidentMaxTargets);
    *ctas = *rows_per_cta = *warps = 0;
    if (!pl.var->small || pl.var->T != 1 || N == 0 || N > max_n || env_int("NBODY_NO_RESIDENT", 0)) return;
    const uint32_t rows = ceil_div(N, 32u);
    const uint32_t per = ceil_div(rows, (uint32_t)kResidentMaxCtas);
    if (per > (uint32_t)kResidentWarps) return; /* thread t owns target t of the CTA's rows */
    if (resident_smem_bytes(pl.src_packs_pad, pl.nslab, per) > kSmallWorldMaxSmem) return;
    const uint32_t units = per * pl.nslab;
    *rows_per_cta = per;
    *ctas = ceil_div(rows, per);
    *warps = units <= 8 && per <= 8 ? 8 : units <= 16 && per <= 16 ? 16 : 32;
}

static void setup_resident(SimPipeline *sim, bool allowed) {
    sim->resident_ctas = 0;
    if (!allowed || sim->allow_persist || sim->total_shards != 1) return;
    resident_geometry(sim->plan, sim->data.total_len, &sim->resident_ctas, &sim->resident_rows, &sim->resident_warps);
    if (!sim->resident_ctas) return;
    const void *func = resident_func(sim->resident_warps);
    NB_CUDA(cudaSetDevice(sim->shards[0].device));
    NB_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallWorldMaxSmem));
    if (sim->resident_ctas > 8) NB_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    resident_launch_config(sim, &cfg, attr, false);
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, func, &cfg) != cudaSuccess || clusters < 1) {
        (void)cudaGetLastError();
        sim->resident_ctas = 0; /* this device cannot co-schedule the cluster: launch per step instead */
    }
}

static SimPipeline *create_pipeline(WorldData data, const NbSimConfig *cfg_in) {
    NbSimConfig cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.device = -1;
    if (cfg_in) {
        NB_ASSERT(cfg_in->struct_size >= 8 && cfg_in->struct_size <= 4096, "NbSimConfig.struct_size not set");
        memcpy(&cfg, cfg_in, std::min<size_t>(cfg_in->struct_size, sizeof cfg));
    }

    NB_ASSERT(data.mass_len <= data.total_len, "mass_len %u > total_len %u", data.mass_len, data.total_len);
    SimPipeline *sim = new SimPipeline();
    sim->data = data;
    sim->data.dt = 0.f;
    if (data.total_len == 0) { /* legal on the reference's CPU path; nothing to allocate, nothing to run */
        sim->inert = true;
        memset(&sim->prop, 0, sizeof sim->prop);
        memset(&sim->plan, 0, sizeof sim->plan);
        return sim;
    }

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        NB_FAIL("no usable CUDA device (%s); this library has no CPU fallback for the GPU path",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    const int first_dev = cfg.device >= 0 ? cfg.device : env_int("NBODY_DEVICE", 0);
    sim->local_gpus = cfg.local_gpus > 0 ? cfg.local_gpus : std::max(1, env_int("NBODY_NGPUS", 1));
    sim->world = std::max(1, cfg.world);
    sim->rank = cfg.world > 1 ? cfg.rank : 0;
    NB_ASSERT(sim->rank >= 0 && sim->rank < sim->world, "rank %d outside world %d", sim->rank, sim->world);
    NB_ASSERT(first_dev >= 0 && first_dev + sim->local_gpus <= ndev, "devices %d..%d requested, %d visible", first_dev,
              first_dev + sim->local_gpus - 1, ndev);
    NB_ASSERT(sim->local_gpus <= kMaxDst, "at most %d GPUs per process", kMaxDst);
    sim->total_shards = sim->world * sim->local_gpus;
    sim->overlap = !(cfg.flags & NB_SIM_FLAG_NO_OVERLAP) && !env_int("NBODY_NO_OVERLAP", 0);
    sim->pdl = !env_int("NBODY_NO_PDL", 0);
    sim->allow_persist = (cfg.flags & NB_SIM_FLAG_PERSIST) || env_int("NBODY_PERSIST", 0);
    sim->use_graph = !(cfg.flags & NB_SIM_FLAG_NO_GRAPH) && !env_int("NBODY_NO_GRAPH", 0);

    NB_CUDA(cudaSetDevice(first_dev));
    NB_CUDA(cudaGetDeviceProperties(&sim->prop, first_dev));
    NB_ASSERT(sim->prop.major >= 10, "device %d (%s) is sm_%d%d; this library is built for sm_100a only", first_dev,
              sim->prop.name, sim->prop.major, sim->prop.minor);
    make_plan(sim, cfg.slab_len ? cfg.slab_len : (uint32_t)env_int("NBODY_SLAB", 0), (cfg.flags & NB_SIM_FLAG_EXACT) != 0);
    const Plan &pl = sim->plan;

    /* who computes which tiles */
    const int S = sim->total_shards;
    sim->all_a0.resize(S); sim->all_a1.resize(S); sim->all_b0.resize(S); sim->all_b1.resize(S);
    for (int q = 0; q < S; q++) shard_tiles(pl, S, q, &sim->all_a0[q], &sim->all_a1[q], &sim->all_b0[q], &sim->all_b1[q]);

    /* exchange mode */
    if (S == 1) {
        sim->exchange = kExchangeNone;
    } else if (sim->world == 1 && !(cfg.flags & NB_SIM_FLAG_NO_PEER_STORE) && !env_int("NBODY_NO_PEER_STORE", 0)) {
        sim->exchange = kExchangePeerStore;
        for (int a = 0; a < sim->local_gpus && sim->exchange == kExchangePeerStore; a++)
            for (int b = 0; b < sim->local_gpus; b++) {
                int ok = 1;
                if (a != b) NB_CUDA(cudaDeviceCanAccessPeer(&ok, first_dev + a, first_dev + b));
                if (!ok) { sim->exchange = kExchangeNccl; break; }
            }
    } else {
        sim->exchange = kExchangeNccl;
    }

    sim->shards.resize(sim->local_gpus);
    for (int l = 0; l < sim->local_gpus; l++) {
        Shard &sh = sim->shards[l];
        sh.device = first_dev + l;
        sh.global_index = sim->rank * sim->local_gpus + l;
        sh.a0 = sim->all_a0[sh.global_index]; sh.a1 = sim->all_a1[sh.global_index];
        sh.b0 = sim->all_b0[sh.global_index]; sh.b1 = sim->all_b1[sh.global_index];
        NB_CUDA(cudaSetDevice(sh.device));
        if (l > 0) {
            NB_CUDA(cudaFuncSetAttribute(plan_func(pl), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan_smem_limit(pl)));
            if (pl.var->cluster && pl.nslab > 8)
                NB_CUDA(cudaFuncSetAttribute(pl.var->func, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        }
        if (cfg.stream && sim->local_gpus == 1) {
            sh.stream = (cudaStream_t)cfg.stream;
            sh.own_stream = false;
        } else {
            NB_CUDA(cudaStreamCreateWithFlags(&sh.stream, cudaStreamNonBlocking));
        }
        NB_CUDA(cudaStreamCreateWithFlags(&sh.comm_stream, cudaStreamNonBlocking));
        NB_CUDA(cudaEventCreate(&sh.ev_begin));
        NB_CUDA(cudaEventCreate(&sh.ev_end));
        NB_CUDA(cudaEventCreateWithFlags(&sh.ev_massive, cudaEventDisableTiming));
        NB_CUDA(cudaEventCreateWithFlags(&sh.ev_comm, cudaEventDisableTiming));
        NB_CUDA(cudaEventCreateWithFlags(&sh.ev_step, cudaEventDisableTiming));
        const size_t N = data.total_len;
        /* one readable failure instead of a bare cudaErrorMemoryAllocation from somewhere in the middle */
        const size_t own_targets = (size_t)((sh.a1 - sh.a0) + (sh.b1 - sh.b0)) * pl.tile;
        const size_t part_bytes = pl.ngroups > 1 ? (size_t)pl.ngroups * own_targets * sizeof(float2) : 0;
        const size_t need = N * (sizeof(Particle) + 3 * sizeof(float2) + sizeof(float)) +
                            (size_t)pl.src_packs_pad * 5 * sizeof(float4) + part_bytes + pl.ntiles * sizeof(unsigned);
        size_t free_b = 0, total_b = 0;
        NB_CUDA(cudaMemGetInfo(&free_b, &total_b));
        NB_ASSERT(need + (64u << 20) <= free_b,
                  "device %d: a World of %u particles (%u massive) needs %.2f GB of device memory (%.2f GB of it slab "
                  "partials: %u per target), %.2f GB are free of %.2f GB",
                  sh.device, data.total_len, data.mass_len, need * 1e-9, part_bytes * 1e-9, pl.ngroups, free_b * 1e-9,
                  total_b * 1e-9);
        sh.aos = dev_alloc<Particle>(N);
        sh.src_xy[0] = dev_alloc<float4>((size_t)pl.src_packs_pad * 2);
        sh.src_xy[1] = dev_alloc<float4>((size_t)pl.src_packs_pad * 2);
        sh.src_gm = dev_alloc<float4>(pl.src_packs_pad);
        sh.pos = dev_alloc<float2>(N);
        sh.vel = dev_alloc<float2>(N);
        sh.acc = dev_alloc<float2>(N);
        sh.radius = dev_alloc<float>(N);
        /* partial sums only for the tiles this GPU computes: [nslab][own tiles * tile] */
        sh.part = dev_alloc<float2>(pl.ngroups > 1 ? (size_t)pl.ngroups * own_targets : 1);
        sh.tile_cnt = dev_alloc<unsigned>(pl.ntiles);
        sh.grid_bar = dev_alloc<unsigned>(1);
        NB_CUDA(cudaMemsetAsync(sh.grid_bar, 0, sizeof(unsigned), sh.stream));
        sh.scratch64 = dev_alloc<double>(8);
        /* on the stream the step kernels run on: the legacy default stream does not order with it */
        NB_CUDA(cudaMemsetAsync(sh.tile_cnt, 0, sizeof(unsigned) * pl.ntiles, sh.stream));
    }
    if (sim->exchange == kExchangePeerStore) {
        for (int a = 0; a < sim->local_gpus; a++) {
            NB_CUDA(cudaSetDevice(first_dev + a));
            for (int b = 0; b < sim->local_gpus; b++) {
                if (a == b) continue;
                cudaError_t pe = cudaDeviceEnablePeerAccess(first_dev + b, 0);
                if (pe == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
                else NB_CUDA(pe);
            }
        }
    }
    if (sim->exchange == kExchangeNccl) {
        NcclApi &api = nccl();
        if (sim->world == 1) {
            std::vector<int> devs(sim->local_gpus);
            std::vector<NcclComm> comms(sim->local_gpus);
            for (int l = 0; l < sim->local_gpus; l++) devs[l] = first_dev + l;
            NB_NCCL(api.CommInitAll(comms.data(), sim->local_gpus, devs.data()));
            for (int l = 0; l < sim->local_gpus; l++) sim->shards[l].comm = comms[l];
        } else {
            NB_ASSERT(cfg.unique_id != nullptr, "world > 1 needs NbSimConfig.unique_id from NbSimGetUniqueId");
            NcclUniqueId id;
            memcpy(&id, cfg.unique_id, sizeof id);
            NB_NCCL(api.GroupStart());
            for (int l = 0; l < sim->local_gpus; l++) {
                NB_CUDA(cudaSetDevice(first_dev + l));
                NB_NCCL(api.CommInitRank(&sim->shards[l].comm, S, id, sim->shards[l].global_index));
            }
            NB_NCCL(api.GroupEnd());
        }
    }
    /* CUDA-graph chunks only where a step is shorter than a launch takes the host (~3.5 us): the one-CTA-per-tile
     * small-World plans on one GPU, on a stream we own (a caller's stream may itself be under capture) */
    sim->use_graph = sim->use_graph && sim->total_shards == 1 && pl.var->small && pl.ntiles <= (uint32_t)sim->prop.multiProcessorCount &&
                     sim->shards[0].own_stream && sim->data.total_len <= kGraphMaxTargets;
    setup_resident(sim, !(cfg.flags & NB_SIM_FLAG_NO_RESIDENT));
    for (Shard &sh : sim->shards) { /* creation-time initialisation done before anything else uses the buffers */
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaStreamSynchronize(sh.stream));
    }
    return sim;
}

static void destroy_pipeline(SimPipeline *sim) {
    if (!sim) return;
    for (StepGraph &g : sim->graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaStreamSynchronize(sh.stream));
        NB_CUDA(cudaStreamSynchronize(sh.comm_stream));
        if (sh.comm) NB_NCCL(nccl().CommDestroy(sh.comm));
        cudaFree(sh.aos); cudaFree(sh.src_xy[0]); cudaFree(sh.src_xy[1]); cudaFree(sh.src_gm);
        cudaFree(sh.pos); cudaFree(sh.vel); cudaFree(sh.acc); cudaFree(sh.radius);
        cudaFree(sh.part); cudaFree(sh.tile_cnt); cudaFree(sh.grid_bar); cudaFree(sh.scratch64);
        cudaEventDestroy(sh.ev_begin); cudaEventDestroy(sh.ev_end); cudaEventDestroy(sh.ev_massive);
        cudaEventDestroy(sh.ev_comm); cudaEventDestroy(sh.ev_step);
        if (sh.own_stream) cudaStreamDestroy(sh.stream);
        cudaStreamDestroy(sh.comm_stream);
    }
    delete sim;
}

}  // namespace nb

/* ---- upload / download ----------------------------------------------------------------- */

namespace nb {

static void sync_all(SimPipeline *sim) {
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaStreamSynchronize(sh.stream));
        NB_CUDA(cudaStreamSynchronize(sh.comm_stream));
    }
}

/* Sources [0, M) of shard q as a pack range (tile is a multiple of 4, so the range starts on a pack). */
static bool source_packs_of_shard(const SimPipeline *sim, int q, size_t *p0, size_t *p1) {
    uint32_t lo, hi;
    particle_range_of_tiles(sim, sim->all_a0[q], sim->all_a1[q], sim->data.mass_len, &lo, &hi);
    *p0 = lo / 4;
    *p1 = ceil_div(hi, 4);
    return hi > lo;
}

/* After an upload every GPU holds the source packs of its own massive particles: replicate them (both position
 * buffers -- the pack that straddles mass_len carries padding lanes no step ever rewrites -- and G*m). */
static void replicate_sources(SimPipeline *sim) {
    if (sim->exchange == kExchangeNccl) {
        NcclApi &api = nccl();
        NB_NCCL(api.GroupStart());
        for (Shard &sh : sim->shards) {
            for (int q = 0; q < sim->total_shards; q++) {
                size_t p0, p1;
                if (!source_packs_of_shard(sim, q, &p0, &p1)) continue;
                for (int b = 0; b < 2; b++) {
                    float *xy = reinterpret_cast<float *>(sh.src_xy[b]) + p0 * 8;
                    NB_NCCL(api.Broadcast(xy, xy, (p1 - p0) * 8, kNcclFloat, q, sh.comm, sh.stream));
                }
                float *gm = reinterpret_cast<float *>(sh.src_gm) + p0 * 4;
                NB_NCCL(api.Broadcast(gm, gm, (p1 - p0) * 4, kNcclFloat, q, sh.comm, sh.stream));
            }
        }
        NB_NCCL(api.GroupEnd());
        return;
    }
    /* one process: the owner pushes its packs to every other GPU once its own conversion is done */
    sync_all(sim);
    for (Shard &src : sim->shards) {
        size_t p0, p1;
        if (!source_packs_of_shard(sim, src.global_index, &p0, &p1)) continue;
        NB_CUDA(cudaSetDevice(src.device));
        for (Shard &dst : sim->shards) {
            if (&dst == &src) continue;
            for (int b = 0; b < 2; b++)
                NB_CUDA(cudaMemcpyPeerAsync(dst.src_xy[b] + p0 * 2, dst.device, src.src_xy[b] + p0 * 2, src.device,
                                            (p1 - p0) * 2 * sizeof(float4), src.stream));
            NB_CUDA(cudaMemcpyPeerAsync(dst.src_gm + p0, dst.device, src.src_gm + p0, src.device,
                                        (p1 - p0) * sizeof(float4), src.stream));
        }
    }
}

/*
 * SetSimulationData.  One GPU: the whole AoS array is copied and converted.  Several GPUs (in this process or
 * across ranks): every GPU copies ONLY the particles it owns from PS (N x 32 B over PCIe in total, not per GPU),
 * converts them, and the source packs (x, y, G*m of the massive particles, 12 B each) are then replicated
 * GPU to GPU -- peer copies inside a process, NCCL broadcasts across processes.  Every rank passes the same
 * array; a rank reads only its own ranges of it.
 */
static void upload(SimPipeline *sim, const Particle *ps) {
    const uint32_t N = sim->data.total_len, M = sim->data.mass_len;
    const Plan &pl = sim->plan;
    const uint32_t data_packs = ceil_div(M, 4u);
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        const uint32_t t0[2] = {sh.a0, sh.b0}, t1[2] = {sh.a1, sh.b1};
        for (int r = 0; r < 2; r++) {
            uint32_t lo = 0, hi = N;
            if (sim->total_shards > 1) particle_range_of_tiles(sim, t0[r], t1[r], N, &lo, &hi);
            else if (r == 1) break;
            if (hi <= lo) continue;
            NB_CUDA(cudaMemcpyAsync(sh.aos + lo, ps + lo, (size_t)(hi - lo) * sizeof(Particle), cudaMemcpyHostToDevice,
                                    sh.stream));
            unpack_targets_kernel<<<ceil_div(hi - lo, 256), 256, 0, sh.stream>>>(sh.aos + lo, hi - lo, sh.pos + lo,
                                                                                  sh.vel + lo, sh.acc + lo, sh.radius + lo);
            sim->kernels_launched++;
        }
        /* own source packs, and the padding packs past mass_len (they need no particle data) */
        size_t p0 = 0, p1 = data_packs;
        const bool own = sim->total_shards == 1 || source_packs_of_shard(sim, sh.global_index, &p0, &p1);
        if (own && p1 > p0) {
            unpack_sources_kernel<<<ceil_div((uint32_t)(p1 - p0), 128), 128, 0, sh.stream>>>(
                sh.aos, M, (uint32_t)p0, (uint32_t)p1, sh.src_xy[0], sh.src_xy[1], sh.src_gm);
            sim->kernels_launched++;
        }
        if (pl.src_packs_pad > data_packs) {
            unpack_sources_kernel<<<ceil_div(pl.src_packs_pad - data_packs, 128), 128, 0, sh.stream>>>(
                sh.aos, M, data_packs, pl.src_packs_pad, sh.src_xy[0], sh.src_xy[1], sh.src_gm);
            sim->kernels_launched++;
        }
        NB_CUDA(cudaGetLastError());
    }
    if (sim->total_shards > 1) replicate_sources(sim);
    sim->cur = 0;
    sim->has_data = true;
    sync_all(sim); /* PS may be pageable and reused by the caller right after we return */
}

/* Patch pos/vel/acc of the ranges a GPU owns into its AoS copy. */
static void pack_own_ranges(SimPipeline *sim, Shard &sh) {
    const uint32_t N = sim->data.total_len;
    uint32_t lo, hi;
    const uint32_t t0[2] = {sh.a0, sh.b0}, t1[2] = {sh.a1, sh.b1};
    for (int r = 0; r < 2; r++) {
        particle_range_of_tiles(sim, t0[r], t1[r], N, &lo, &hi);
        if (hi <= lo) continue;
        pack_targets_kernel<<<ceil_div(hi - lo, 256), 256, 0, sh.stream>>>(sh.aos, lo, hi, sh.pos, sh.vel, sh.acc);
        NB_CUDA(cudaGetLastError());
        sim->kernels_launched++;
    }
}

/* LOCAL_ONLY: only the ranges owned by this process's GPUs are written to PS (no gather across ranks). */
static void download(SimPipeline *sim, Particle *ps, bool local_only = false) {
    const uint32_t N = sim->data.total_len;
    NB_ASSERT(sim->has_data, "GetSimulationData before SetSimulationData");
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        pack_own_ranges(sim, sh);
    }
    if (sim->world > 1 && !local_only) {
        /* other processes own some ranges: all-gather the AoS ranges (32 B = 8 floats per particle) */
        NcclApi &api = nccl();
        NB_NCCL(api.GroupStart());
        for (Shard &sh : sim->shards) {
            for (int q = 0; q < sim->total_shards; q++) {
                uint32_t lo, hi;
                const uint32_t t0[2] = {sim->all_a0[q], sim->all_b0[q]}, t1[2] = {sim->all_a1[q], sim->all_b1[q]};
                for (int r = 0; r < 2; r++) {
                    particle_range_of_tiles(sim, t0[r], t1[r], N, &lo, &hi);
                    if (hi <= lo) continue;
                    NB_NCCL(api.Broadcast(sh.aos + lo, sh.aos + lo, (size_t)(hi - lo) * 8, kNcclFloat, q, sh.comm, sh.stream));
                }
            }
        }
        NB_NCCL(api.GroupEnd());
        Shard &sh = sim->shards[0];
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaMemcpyAsync(ps, sh.aos, (size_t)N * sizeof(Particle), cudaMemcpyDeviceToHost, sh.stream));
    } else {
        for (Shard &sh : sim->shards) {
            NB_CUDA(cudaSetDevice(sh.device));
            uint32_t lo, hi;
            const uint32_t t0[2] = {sh.a0, sh.b0}, t1[2] = {sh.a1, sh.b1};
            for (int r = 0; r < 2; r++) {
                particle_range_of_tiles(sim, t0[r], t1[r], N, &lo, &hi);
                if (hi <= lo) continue;
                NB_CUDA(cudaMemcpyAsync(ps + lo, sh.aos + lo, (size_t)(hi - lo) * sizeof(Particle),
                                        cudaMemcpyDeviceToHost, sh.stream));
            }
        }
    }
    sync_all(sim);
}

}  // namespace nb

/* ---- stepping ------------------------------------------------------------------------------ */

namespace nb {

static void fill_params(SimPipeline *sim, int l, float dt, StepParams *p) {
    const Plan &pl = sim->plan;
    Shard &sh = sim->shards[l];
    const int cur = sim->cur, nxt = cur ^ 1;
    memset(p, 0, sizeof *p);
    p->src_xy = sh.src_xy[cur];
    p->src_gm = sh.src_gm;
    p->dst_xy[0] = reinterpret_cast<float *>(sh.src_xy[nxt]);
    p->ndst = 1;
    if (sim->exchange == kExchangePeerStore) {
        for (int o = 0; o < sim->local_gpus; o++)
            if (o != l) p->dst_xy[p->ndst++] = reinterpret_cast<float *>(sim->shards[o].src_xy[nxt]);
    }
    p->pos = sh.pos; p->vel = sh.vel; p->acc = sh.acc; p->radius = sh.radius;
    p->part = sh.part; p->tile_cnt = sh.tile_cnt;
    p->total_len = sim->data.total_len; p->mass_len = sim->data.mass_len;
    p->part_stride = ((sh.a1 - sh.a0) + (sh.b1 - sh.b0)) * pl.tile;
    p->src_packs_pad = pl.src_packs_pad;
    /* the kernel's "slab" is what one CTA walks: a slab, or a group of slabs it folds itself (GROUPED) */
    p->slab_packs = pl.slab_packs * pl.group_slabs;
    p->nslab = pl.var->small ? pl.nslab : pl.ngroups; /* small Worlds: all slabs inside the tile's one CTA */
    p->fold_tiles = pl.group_slabs > 1 ? pl.slab_packs / (uint32_t)pl.var->tjp : 0;
    p->dt = dt;
}

/* Point P at the target tiles [t0, t1) of this GPU; returns the grid size. */
static uint32_t aim_at_tiles(SimPipeline *sim, Shard &sh, StepParams &p, uint32_t t0, uint32_t t1) {
    p.tile_first = t0;
    /* partial slots are local: the massive tiles [a0, a1) come first, the massless tiles [b0, b1) follow.
     * The kernel indexes by global particle number, so the base pointer is biased here (never dereferenced
     * outside the allocation); an extra kernel parameter for this cost 2.4 % at N = 1M through a different
     * ptxas schedule of the inner loop. */
    const size_t first = (size_t)(t0 >= sh.b0 && t0 >= sh.a1 ? sh.b0 - (sh.a1 - sh.a0) : sh.a0) * sim->plan.tile;
    p.part = sh.part ? reinterpret_cast<float2 *>(reinterpret_cast<uintptr_t>(sh.part) - first * sizeof(float2)) : nullptr;
    return (t1 - t0) * sim->plan.ngroups;
}

static void launch_tiles(SimPipeline *sim, Shard &sh, StepParams &p, uint32_t t0, uint32_t t1) {
    if (t1 <= t0) return;
    uint32_t grid = aim_at_tiles(sim, sh, p, t0, t1);
    if (sim->plan.var->balanced) { /* the units of rows [t0, t1) over at most one CTA per SM */
        p.fold_tiles = t1 - t0;
        grid = std::min<uint32_t>(grid, (uint32_t)sim->prop.multiProcessorCount);
    }
    /* Programmatic dependent launch on every launch: the kernel releases its dependents after its pair
     * loop, so the next step's CTAs are scheduled into slots as they drain and have their prologue done when
     * the last partial lands.  (Released at kernel entry instead, the next step's CTAs were parked early on an
     * uneven set of SMs: N = 4 000 went from 10.3 to 13.7 us per step; released late, 10.3 -> 9.2.) */
    const bool pdl = sim->pdl;
    if (sim->plan.csplit > 1) launch_small_split<32, 1, 2>(grid * sim->plan.csplit, sim->plan.csplit, sh.stream, p, pdl);
    else if (sim->plan.group_slabs > 1) sim->plan.var->launch_grouped(grid, sh.stream, p, pdl);
    else sim->plan.var->launch(grid, sh.stream, p, pdl);
    NB_CUDA(cudaGetLastError());
    sim->kernels_launched++;
}

/* Re-replicate the new massive positions (buffer `nxt`) with NCCL: every shard broadcasts its range. */
static void exchange_nccl(SimPipeline *sim, int nxt, bool on_comm_stream) {
    NcclApi &api = nccl();
    const uint32_t M = sim->data.mass_len;
    NB_NCCL(api.GroupStart());
    for (Shard &sh : sim->shards) {
        float *buf = reinterpret_cast<float *>(sh.src_xy[nxt]);
        for (int q = 0; q < sim->total_shards; q++) {
            uint32_t lo, hi;
            particle_range_of_tiles(sim, sim->all_a0[q], sim->all_a1[q], M, &lo, &hi);
            if (hi <= lo) continue;
            const size_t p0 = lo / 4, p1 = ceil_div(hi, 4); /* tile is a multiple of 4: lo is pack aligned */
            NB_NCCL(api.Broadcast(buf + p0 * 8, buf + p0 * 8, (p1 - p0) * 8, kNcclFloat, q, sh.comm,
                                  on_comm_stream ? sh.comm_stream : sh.stream));
        }
    }
    NB_NCCL(api.GroupEnd());
}

/*
 * Chained steps of a small World as CUDA-graph chunks.  With the small-World kernel a step at N <= ~1 500 takes the
 * GPU less time than one cudaLaunchKernelEx takes the host (tools/launch_bound_probe.py: 3.5 us per launch in a
 * 100-step call, device time 3.5 us per step = launch-bound).  A chunk of kGraphSteps launches -- the same kernels,
 * arguments and programmatic-dependent-launch edges as the stream path, captured once per (dt, starting buffer) --
 * is replayed with one cudaGraphLaunch; an even chunk length brings the ping-pong buffers back to where they
 * started, so the same graph serves every chunk of a call.  The first chained call (n >= 2) prepares the graph.
 */
static void launch_tiles(SimPipeline *sim, Shard &sh, StepParams &p, uint32_t t0, uint32_t t1);
static void fill_params(SimPipeline *sim, int l, float dt, StepParams *p);

static cudaGraphExec_t step_graph(SimPipeline *sim, float dt) {
    StepGraph &g = sim->graphs[sim->cur];
    if (g.exec && g.dt == dt) return g.exec;
    if (g.exec) {
        NB_CUDA(cudaGraphExecDestroy(g.exec));
        g.exec = nullptr;
    }
    Shard &sh = sim->shards[0];
    const int cur0 = sim->cur;
    const uint64_t launched0 = sim->kernels_launched;
    NB_CUDA(cudaStreamBeginCapture(sh.stream, cudaStreamCaptureModeThreadLocal));
    for (uint32_t k = 0; k < kGraphSteps; k++) {
        StepParams p;
        fill_params(sim, 0, dt, &p);
        launch_tiles(sim, sh, p, sh.a0, sh.b1);
        sim->cur ^= 1;
    }
    cudaGraph_t graph = nullptr;
    NB_CUDA(cudaStreamEndCapture(sh.stream, &graph));
    sim->cur = cur0;
    sim->kernels_launched = launched0; /* recorded, not launched */
    NB_CUDA(cudaGraphInstantiate(&g.exec, graph, 0));
    NB_CUDA(cudaGraphDestroy(graph));
    g.dt = dt;
    return g.exec;
}

static void step_async(SimPipeline *sim, uint32_t n, float dt) {
    NB_ASSERT(sim->has_data, "PerformSimUpdate before SetSimulationData (ref sim_gpu.h:34)");
    const int L = sim->local_gpus;
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaEventRecord(sh.ev_begin, sh.stream));
    }
    const bool split = sim->exchange == kExchangeNccl && sim->overlap;
    if (sim->resident_ctas && n >= 2 && sim->shards[0].a1 == sim->shards[0].b0) {
        /* a tiny World: all n steps in one launch of one cluster, state resident on chip (resident_kernel.cuh) */
        Shard &sh = sim->shards[0];
        NB_CUDA(cudaSetDevice(sh.device));
        ResidentParams rp;
        fill_params(sim, 0, dt, &rp.s);
        rp.s.dst_xy[0] = reinterpret_cast<float *>(sh.src_xy[(sim->cur + (int)(n & 1u)) & 1]); /* where the last step's positions go */
        rp.src_xy_alt = sh.src_xy[sim->cur ^ 1];
        rp.nsteps = n;
        rp.rows_per_cta = sim->resident_rows;
        cudaLaunchConfig_t cfg;
        cudaLaunchAttribute attr[2];
        resident_launch_config(sim, &cfg, attr, sim->pdl);
        void *args[1] = {&rp};
        NB_CUDA(cudaLaunchKernelExC(&cfg, resident_func(sim->resident_warps), args));
        sim->kernels_launched++;
        sim->resident_launches++;
        sim->cur ^= (int)(n & 1u);
        n = 0;
    }
    if (sim->use_graph && n >= 2 && sim->shards[0].a1 == sim->shards[0].b0) {
        static_assert(kGraphSteps % 2 == 0, "a chunk must bring the ping-pong buffers back");
        Shard &sh = sim->shards[0];
        cudaGraphExec_t exec = step_graph(sim, dt);
        while (n >= kGraphSteps) {
            NB_CUDA(cudaGraphLaunch(exec, sh.stream));
            sim->kernels_launched += kGraphSteps;
            sim->graph_launches++;
            n -= kGraphSteps;
        }
    }
    /* Opt-in (NBODY_PERSIST=1): small World, several steps: one cooperative launch of the persistent kernel runs all
     * of them, a grid barrier between steps instead of a kernel boundary (persist_kernel.cuh).  Same plan, same
     * bits -- and measured slower than the launches below, which is why it is not the default. */
    if (n >= 2 && L == 1 && sim->exchange == kExchangeNone && sim->persist_ctas > 0) {
        Shard &sh = sim->shards[0];
        PersistParams pp;
        fill_params(sim, 0, dt, &pp.s);
        const uint32_t grid = aim_at_tiles(sim, sh, pp.s, sh.a0, sh.b1);
        if (sh.a1 == sh.b0 && grid > 0 && grid <= sim->persist_ctas) {
            pp.bar = sh.grid_bar;
            pp.bar_base = sh.bar_base;
            pp.nsteps = n;
            sim->plan.var->launch_persist(grid, sh.stream, pp);
            NB_CUDA(cudaGetLastError());
            sh.bar_base += (n - 1) * grid; /* the last step ends at the kernel boundary, not at the barrier */
            sim->kernels_launched++;
            sim->persistent_launches++;
            sim->cur ^= (int)(n & 1u);
            n = 0; /* nothing left for the launch-per-step loop */
        }
    }
    for (uint32_t it = 0; it < n; it++) {
        const int nxt = sim->cur ^ 1;
        for (int l = 0; l < L; l++) {
            Shard &sh = sim->shards[l];
            NB_CUDA(cudaSetDevice(sh.device));
            StepParams p;
            fill_params(sim, l, dt, &p);
            if (sh.a1 == sh.b0 && !split) {
                launch_tiles(sim, sh, p, sh.a0, sh.b1); /* one contiguous launch */
            } else {
                launch_tiles(sim, sh, p, sh.a0, sh.a1); /* massive targets first: they feed the exchange */
                if (split) {
                    NB_CUDA(cudaEventRecord(sh.ev_massive, sh.stream));
                    NB_CUDA(cudaStreamWaitEvent(sh.comm_stream, sh.ev_massive, 0));
                }
                launch_tiles(sim, sh, p, sh.b0, sh.b1); /* massless targets overlap the exchange */
            }
        }
        if (sim->exchange == kExchangeNccl) {
            exchange_nccl(sim, nxt, split);
            if (split) {
                for (Shard &sh : sim->shards) {
                    NB_CUDA(cudaSetDevice(sh.device));
                    NB_CUDA(cudaEventRecord(sh.ev_comm, sh.comm_stream));
                    NB_CUDA(cudaStreamWaitEvent(sh.stream, sh.ev_comm, 0));
                }
            }
        } else if (sim->exchange == kExchangePeerStore) {
            /* positions were stored into every peer by the epilogue; the next step of each GPU
             * must wait until all GPUs finished this one */
            for (Shard &sh : sim->shards) {
                NB_CUDA(cudaSetDevice(sh.device));
                NB_CUDA(cudaEventRecord(sh.ev_step, sh.stream));
            }
            for (int l = 0; l < L; l++) {
                NB_CUDA(cudaSetDevice(sim->shards[l].device));
                for (int o = 0; o < L; o++)
                    if (o != l) NB_CUDA(cudaStreamWaitEvent(sim->shards[l].stream, sim->shards[o].ev_step, 0));
            }
        }
        sim->cur = nxt;
    }
    for (Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaEventRecord(sh.ev_end, sh.stream));
    }
    sim->timed = true;
}

}  // namespace nb

/* ---- C ABI (include/nbody_sim.h) -------------------------------------------------------- */

extern "C" {

SimPipeline *CreateSimPipeline(WorldData data) {
    nb::DeviceGuard guard;
    return nb::create_pipeline(data, nullptr);
}

SimPipeline *NbSimCreate(WorldData data, const NbSimConfig *cfg) {
    nb::DeviceGuard guard;
    return nb::create_pipeline(data, cfg);
}

void DestroySimPipeline(SimPipeline *sim) {
    nb::DeviceGuard guard;
    nb::destroy_pipeline(sim);
}

void SetSimulationData(SimPipeline *sim, const Particle *ps) {
    if (sim->inert) return;
    nb::DeviceGuard guard;
    nb::upload(sim, ps);
}

void GetSimulationData(const SimPipeline *sim, Particle *ps) {
    if (sim->inert) return;
    nb::DeviceGuard guard;
    nb::download(const_cast<SimPipeline *>(sim), ps);
}

void NbSimGetDataLocal(SimPipeline *sim, Particle *ps) {
    if (sim->inert) return;
    nb::DeviceGuard guard;
    nb::download(sim, ps, true);
}

uint32_t NbSimLocalRanges(const SimPipeline *sim, uint32_t *out, uint32_t max_ranges) {
    uint32_t n = 0;
    for (const nb::Shard &sh : sim->shards) {
        const uint32_t t0[2] = {sh.a0, sh.b0}, t1[2] = {sh.a1, sh.b1};
        for (int r = 0; r < 2; r++) {
            uint32_t lo, hi;
            nb::particle_range_of_tiles(sim, t0[r], t1[r], sim->data.total_len, &lo, &hi);
            if (hi <= lo) continue;
            if (n < max_ranges) {
                out[2 * n] = lo;
                out[2 * n + 1] = hi;
            }
            n++;
        }
    }
    return n;
}

void PerformSimUpdate(SimPipeline *sim, uint32_t n, float dt) {
    if (n == 0 || sim->inert) return;
    nb::DeviceGuard guard;
    nb::step_async(sim, n, dt);
    nb::sync_all(sim); /* synchronous like the reference's fence wait, ref sim_gpu.c:352-353 */
}

void NbSimStepAsync(SimPipeline *sim, uint32_t n, float dt) {
    if (n == 0 || sim->inert) return;
    nb::DeviceGuard guard;
    nb::step_async(sim, n, dt);
}

void NbSimSynchronize(SimPipeline *sim) {
    if (sim->inert) return;
    nb::DeviceGuard guard;
    nb::sync_all(sim);
}

float NbSimLastStepMs(SimPipeline *sim) {
    if (!sim->timed) return 0.f;
    nb::DeviceGuard guard;
    float worst = 0.f;
    for (nb::Shard &sh : sim->shards) {
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaEventSynchronize(sh.ev_end));
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, sh.ev_begin, sh.ev_end));
        worst = ms > worst ? ms : worst;
    }
    return worst;
}

void NbSimGetUniqueId(void *out) {
    nb::NcclUniqueId id;
    NB_NCCL(nb::nccl().GetUniqueId(&id));
    memcpy(out, &id, NB_SIM_UNIQUE_ID_BYTES);
}

void *NbSimHostAlloc(size_t bytes) {
    void *p = nullptr;
    NB_CUDA(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable));
    return p;
}

void NbSimHostFree(void *p) {
    if (p) NB_CUDA(cudaFreeHost(p));
}

void NbSimPlan(uint32_t total_len, uint32_t mass_len, int32_t shards, int32_t sm_count, NbSimPlanInfo *out,
               uint32_t *tile_ranges) {
    NB_ASSERT(mass_len <= total_len && total_len > 0 && shards >= 1, "bad plan request");
    nb::Plan pl;
    nb::plan_host(total_len, mass_len, sm_count > 0 ? sm_count : 148, (uint32_t)nb::env_int("NBODY_SLAB", 0), pl);
    memset(out, 0, sizeof *out);
    strncpy(out->variant, pl.var->name, sizeof out->variant - 1);
    out->threads = pl.var->threads;
    out->tile_targets = pl.tile;
    out->tiles = pl.ntiles;
    out->massive_tiles = pl.mtiles;
    out->slab_len = pl.slab_packs * 4;
    out->slabs = pl.nslab;
    out->ctas_per_sm = pl.ctas_per_sm;
    out->group_slabs = pl.group_slabs;
    out->groups = pl.var->small ? pl.nslab : pl.ngroups; /* small Worlds: one sum per slab, in shared memory */
    out->cluster_split = pl.csplit;
    if (shards == 1) nb::resident_geometry(pl, total_len, &out->resident_ctas, &out->resident_rows, &out->resident_warps);
    if (tile_ranges)
        for (int q = 0; q < shards; q++)
            nb::shard_tiles(pl, shards, q, &tile_ranges[4 * q], &tile_ranges[4 * q + 1], &tile_ranges[4 * q + 2],
                            &tile_ranges[4 * q + 3]);
}

int32_t NbSimDeviceCount(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return n;
}

void NbSimGetInfo(const SimPipeline *sim, NbSimInfo *out) {
    memset(out, 0, sizeof *out);
    if (sim->inert) { /* an empty World: nothing on any device */
        strncpy(out->device_name, "(none: empty World)", sizeof out->device_name - 1);
        out->device = -1;
        out->world = 1;
        out->kernels_launched = 0;
        return;
    }
    strncpy(out->device_name, sim->prop.name, sizeof out->device_name - 1);
    out->device = sim->shards[0].device;
    out->local_gpus = sim->local_gpus;
    out->rank = sim->rank;
    out->world = sim->world;
    out->sm_count = sim->prop.multiProcessorCount;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, sim->shards[0].device);
    out->sm_clock_khz = khz;
    out->cc_major = sim->prop.major;
    out->cc_minor = sim->prop.minor;
    out->total_len = sim->data.total_len;
    out->mass_len = sim->data.mass_len;
    out->tile_targets = sim->plan.tile;
    out->threads = sim->plan.var->threads;
    out->slab_len = sim->plan.slab_packs * 4;
    out->slabs = sim->plan.nslab;
    out->group_slabs = sim->plan.group_slabs;
    out->persistent_ctas = sim->persist_ctas;
    out->persistent_launches = sim->persistent_launches;
    strncpy(out->variant, sim->plan.var->name, sizeof out->variant - 1);
    out->graph_launches = sim->graph_launches;
    out->graph_steps = sim->use_graph ? nb::kGraphSteps : 0;
    out->cluster_split = sim->plan.csplit;
    out->resident_ctas = sim->resident_ctas;
    out->resident_launches = sim->resident_launches;
    uint32_t tiles = 0;
    for (const nb::Shard &sh : sim->shards) tiles += (sh.a1 - sh.a0) + (sh.b1 - sh.b0);
    out->tiles_local = tiles;
    out->kernels_launched = sim->kernels_launched;
    out->pairs_per_step = (uint64_t)sim->data.total_len * sim->data.mass_len;
}

/* Bring the AoS copy on the first GPU up to date with the whole current state (collective when
 * several processes share the World: every rank must call it). */
static void refresh_aos_on_first_gpu(SimPipeline *sim) {
    const uint32_t N = sim->data.total_len;
    nb::Shard &sh = sim->shards[0];
    NB_ASSERT(sim->has_data, "no particle data on the device yet");
    NB_CUDA(cudaSetDevice(sh.device));
    if (sim->total_shards > 1) {
        Particle *tmp = (Particle *)NbSimHostAlloc((size_t)N * sizeof(Particle));
        nb::download(sim, tmp);
        NB_CUDA(cudaSetDevice(sh.device));
        NB_CUDA(cudaMemcpyAsync(sh.aos, tmp, (size_t)N * sizeof(Particle), cudaMemcpyHostToDevice, sh.stream));
        NB_CUDA(cudaStreamSynchronize(sh.stream));
        NbSimHostFree(tmp);
    } else {
        NB_CUDA(cudaSetDevice(sh.device));
        nb::pack_own_ranges(sim, sh);
    }
}

void NbSimDiagnostics(SimPipeline *sim, double *out7) {
    if (sim->inert) {
        memset(out7, 0, 7 * sizeof(double));
        return;
    }
    nb::DeviceGuard guard;
    const uint32_t N = sim->data.total_len;
    nb::Shard &sh = sim->shards[0];
    refresh_aos_on_first_gpu(sim);
    NB_CUDA(cudaSetDevice(sh.device));
    NB_CUDA(cudaMemsetAsync(sh.scratch64, 0, 7 * sizeof(double), sh.stream));
    nb::diagnostics_kernel<<<nb::ceil_div(N, 256), 256, 0, sh.stream>>>(sh.aos, N, sim->data.mass_len, sh.scratch64);
    NB_CUDA(cudaGetLastError());
    sim->kernels_launched++;
    NB_CUDA(cudaMemcpyAsync(out7, sh.scratch64, 7 * sizeof(double), cudaMemcpyDeviceToHost, sh.stream));
    NB_CUDA(cudaStreamSynchronize(sh.stream));
}

void NbSimFitCamera(SimPipeline *sim, int32_t width, int32_t height, float *target_x, float *target_y, float *zoom) {
    if (sim->inert) {
        *target_x = *target_y = 0.f;
        *zoom = 1.f;
        return;
    }
    nb::DeviceGuard guard;
    const uint32_t N = sim->data.total_len;
    nb::Shard &sh = sim->shards[0];
    refresh_aos_on_first_gpu(sim);
    NB_CUDA(cudaSetDevice(sh.device));
    int *d = nb::dev_alloc<int>(4);
    const int init[4] = {nb::float_to_ordered_host(INFINITY), nb::float_to_ordered_host(INFINITY),
                         nb::float_to_ordered_host(-INFINITY), nb::float_to_ordered_host(-INFINITY)};
    NB_CUDA(cudaMemcpyAsync(d, init, sizeof init, cudaMemcpyHostToDevice, sh.stream));
    nb::bounds_kernel<<<nb::ceil_div(N, 256), 256, 0, sh.stream>>>(sh.aos, N, d);
    NB_CUDA(cudaGetLastError());
    sim->kernels_launched++;
    int h[4];
    NB_CUDA(cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, sh.stream));
    NB_CUDA(cudaStreamSynchronize(sh.stream));
    NB_CUDA(cudaFree(d));
    const float minx = nb::ordered_to_float_host(h[0]), miny = nb::ordered_to_float_host(h[1]);
    const float maxx = nb::ordered_to_float_host(h[2]), maxy = nb::ordered_to_float_host(h[3]);
    /* ref src/main.c:212-222: fit 90 % of the window, never zoom in beyond 1 */
    const float zx = 0.9f * (float)width / (maxx - minx), zy = 0.9f * (float)height / (maxy - miny);
    *zoom = (zx < 1.f || zy < 1.f) ? fminf(zx, zy) : 1.f;
    *target_x = (minx + maxx) * 0.5f;
    *target_y = (miny + maxy) * 0.5f;
}

void NbSimRenderRGBA(SimPipeline *sim, int32_t width, int32_t height, float target_x, float target_y, float zoom,
                     float min_radius, uint8_t *rgba) {
    NB_ASSERT(width > 0 && height > 0 && (int64_t)width * height <= (1 << 28), "bad image size %d x %d", width, height);
    const uint32_t N = sim->data.total_len;
    const uint32_t pixels = (uint32_t)width * (uint32_t)height;
    if (sim->inert) { /* background only */
        for (uint32_t k = 0; k < pixels; k++) rgba[4 * k] = rgba[4 * k + 1] = rgba[4 * k + 2] = 22, rgba[4 * k + 3] = 255;
        return;
    }
    nb::DeviceGuard guard;
    nb::Shard &sh = sim->shards[0];
    refresh_aos_on_first_gpu(sim);
    NB_CUDA(cudaSetDevice(sh.device));
    unsigned *winner = nb::dev_alloc<unsigned>(pixels);
    uchar4 *img = nb::dev_alloc<uchar4>(pixels);
    NB_CUDA(cudaMemsetAsync(winner, 0, (size_t)pixels * sizeof(unsigned), sh.stream));
    const nb::Camera cam = {target_x, target_y, zoom, width, height};
    nb::splat_kernel<<<nb::ceil_div(N, 8), 256, 0, sh.stream>>>(sh.aos, N, cam, min_radius, winner);
    nb::resolve_kernel<<<nb::ceil_div(pixels, 256), 256, 0, sh.stream>>>(sh.aos, winner, pixels, img);
    NB_CUDA(cudaGetLastError());
    sim->kernels_launched += 2;
    NB_CUDA(cudaMemcpyAsync(rgba, img, (size_t)pixels * 4, cudaMemcpyDeviceToHost, sh.stream));
    NB_CUDA(cudaStreamSynchronize(sh.stream));
    NB_CUDA(cudaFree(winner));
    NB_CUDA(cudaFree(img));
}

uint32_t NbSimDebugTimeline(uint64_t *out, uint32_t max_warps) {
    const uint32_t n = std::min<uint32_t>(max_warps, nb::kTimelineWarps);
    NB_CUDA(cudaMemcpyFromSymbol(out, nb::nb_timeline, (size_t)n * 4 * sizeof(uint64_t)));
    return n;
}

void NbSimAccF64(SimPipeline *sim, uint32_t first, uint32_t last, double *out) {
    NB_ASSERT(first <= last && last <= sim->data.total_len, "bad target range [%u, %u)", first, last);
    NB_ASSERT(sim->total_shards == 1, "NbSimAccF64 is a single-GPU diagnostic");
    if (first == last) return;
    nb::DeviceGuard guard;
    nb::Shard &sh = sim->shards[0];
    NB_CUDA(cudaSetDevice(sh.device));
    double *d = nb::dev_alloc<double>(2 * (size_t)(last - first));
    nb::acc_f64_kernel<<<nb::ceil_div(last - first, 128), 128, 0, sh.stream>>>(
        sh.pos, sh.radius, sh.src_xy[sim->cur], sh.src_gm, sim->data.mass_len, first, last, d);
    NB_CUDA(cudaGetLastError());
    sim->kernels_launched++;
    NB_CUDA(cudaMemcpyAsync(out, d, 2 * (size_t)(last - first) * sizeof(double), cudaMemcpyDeviceToHost, sh.stream));
    NB_CUDA(cudaStreamSynchronize(sh.stream));
    NB_CUDA(cudaFree(d));
}

} /* extern "C" */
