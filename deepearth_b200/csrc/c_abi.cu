// c_abi.cu -- the extern "C" surface of libde_grid4d.so (declared in include/de_grid4d.h).
// Host-only logic: argument validation, segmentation of the level list into launches, group / tile
// geometry.  No torch types, no device allocation, no global mutable state (the last-error text is
// thread local), nothing synchronises.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/de_grid4d.h"
#include "grid4d_common.cuh"
#include "launch.h"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
}

bool is_pow2(uint32_t v) { return v && !(v & (v - 1)); }

// L2 working-set budget of one level group (B200 L2 ~126 MB, split over two dies)
constexpr unsigned long long kL2BudgetBytes = 64ull << 20;

int floor_pow2(int v) {
    int p = 1;
    while (p * 2 <= v) p *= 2;
    return p;
}

struct CallArgs {
    const de_level_desc* levels;
    int n_levels;
    const float* coords;
    int64_t n_points;
    int coord_stride;
    int table_dtype;
    int n_features;
    uint32_t flags;
    int levels_per_group;
    bool need_tables, need_grads;
    void* workspace = nullptr;
    size_t workspace_bytes = 0;
};

int validate(const CallArgs& a) {
    if (!a.levels || a.n_levels <= 0 || a.n_levels > DE_MAX_LEVELS_PER_CALL)
        return fail(DE_EINVAL, "n_levels=%d outside 1..%d", a.n_levels, DE_MAX_LEVELS_PER_CALL);
    if (a.n_points < 0) return fail(DE_EINVAL, "n_points=%lld < 0", (long long)a.n_points);
    if (a.n_points > 0 && !a.coords) return fail(DE_EINVAL, "coords is NULL");
    if (a.coord_stride < 1) return fail(DE_EINVAL, "coord_stride=%d < 1", a.coord_stride);
    if (a.n_features < 1 || a.n_features > 64) return fail(DE_EUNSUPPORTED, "n_features=%d outside 1..64", a.n_features);
    if (a.table_dtype != DE_F32 && a.table_dtype != DE_F16) return fail(DE_EUNSUPPORTED, "table_dtype=%d", a.table_dtype);
    if (a.flags & ~DE_FLAG_MASK) return fail(DE_EINVAL, "unknown flags 0x%x", a.flags);
    if (a.levels_per_group < 0 || a.levels_per_group > de::kMaxGroup)
        return fail(DE_EINVAL, "levels_per_group=%d outside 0..%d", a.levels_per_group, de::kMaxGroup);
    for (int i = 0; i < a.n_levels; ++i) {
        const de_level_desc& L = a.levels[i];
        if (L.n_dims < 1 || L.n_dims > DE_MAX_DIMS) return fail(DE_EUNSUPPORTED, "level %d: n_dims=%d outside 1..4", i, L.n_dims);
        if (L.table_size == 0) return fail(DE_EINVAL, "level %d: table_size=0", i);
        if (L.resolution > (1u << 30)) return fail(DE_EUNSUPPORTED, "level %d: resolution %u > 2^30", i, L.resolution);
        for (int d = 0; d < L.n_dims; ++d)
            if (L.dim[d] >= a.coord_stride) return fail(DE_EINVAL, "level %d: dim[%d]=%d >= coord_stride=%d", i, d, L.dim[d], a.coord_stride);
        if (L.dense) {
            unsigned long long cells = 1;
            for (int d = 0; d < L.n_dims; ++d) cells *= (unsigned long long)L.resolution + 1ull;
            if (cells > L.table_size) return fail(DE_EINVAL, "level %d: dense index needs (R+1)^D=%llu <= T=%u", i, cells, L.table_size);
        }
        if (L.mask_id > 2) return fail(DE_EINVAL, "level %d: mask_id=%d", i, L.mask_id);
        if (a.need_tables && !L.table) return fail(DE_EINVAL, "level %d: table is NULL", i);
        if (a.need_grads && !L.grad_table) return fail(DE_EINVAL, "level %d: grad_table is NULL", i);
        const size_t entry = (size_t)a.n_features * (a.table_dtype == DE_F16 ? 2 : 4);
        const size_t align = (a.n_features == 1 || a.n_features == 2 || a.n_features == 4 || a.n_features == 8) ? (entry > 16 ? 16 : entry) : 1;
        if (a.need_tables && (reinterpret_cast<uintptr_t>(L.table) % align)) return fail(DE_EALIGN, "level %d: table not %zu-byte aligned", i, align);
        const size_t galign = (a.n_features == 2) ? 8 : ((a.n_features == 4 || a.n_features == 8) ? 16 : 4);
        if (a.need_grads && (reinterpret_cast<uintptr_t>(L.grad_table) % galign)) return fail(DE_EALIGN, "level %d: grad_table not %zu-byte aligned", i, galign);
    }
    return DE_OK;
}

void fill_level(const de_level_desc& L, de::LevelParams& o) {
    o.table = L.table;
    o.grad = L.grad_table;
    o.T = L.table_size;
    o.res_f = static_cast<float>(L.resolution);   // Python int -> fp32 scalar, round to nearest
    o.res_p1 = L.resolution + 1u;
    o.out_col = L.out_col;
    o.pad0 = 0;
    for (int d = 0; d < 4; ++d) o.dim[d] = L.dim[d];
    o.dense = L.dense ? 1 : 0;
    o.pow2 = is_pow2(L.table_size) ? 1 : 0;
    o.mask_id = L.mask_id;
    o.pad = 0;
    o.rep_base = nullptr;
    o.rep_count = 0;
    o.rep_stride = 0;
    o.rep_live = 0;
    o.pad2 = 0;
    static const uint32_t kPrimes[4] = {1u, 2654435761u, 805459861u, 3674653429u};   // hash_encoding.py:64
    uint64_t stride = 1;
    for (int d = 0; d < 4; ++d) {
        o.mul[d] = L.dense ? static_cast<uint32_t>(stride) : kPrimes[d];
        stride *= static_cast<uint64_t>(L.resolution) + 1ull;   // validate() bounds (R+1)^D <= T < 2^32 for dense
        if (stride > 0xffffffffull) stride = 0;                  // never used: beyond n_dims of a valid dense level
    }
}

bool rows_vectorisable(const void* base, int64_t stride, const de_level_desc* levels, int n_levels, int F) {
    if (!(F == 2 || F == 4 || F == 8)) return false;
    const int unit = F == 2 ? 2 : 4;   // floats per vector access
    if (reinterpret_cast<uintptr_t>(base) % (unit * 4)) return false;
    if (stride % unit) return false;
    for (int i = 0; i < n_levels; ++i)
        if (levels[i].out_col % unit) return false;
    return true;
}

enum class Op { Forward, Backward, Indices };

// The TMA-staged backward needs float4 coordinate rows, every group's columns contiguous in the gradient
// row, and each group's slice of every row 16-byte aligned with a multiple-of-16 length.
bool staging_ok(const CallArgs& a, const de::EncodeParams& P, const int* idx, int n) {
    if (!(a.flags & DE_FLAG_STAGED_BWD)) return false;
    if (!P.vec_coords) return false;
    if (reinterpret_cast<uintptr_t>(P.grad_out) % 16 || P.out_stride % 4) return false;
    const int G = P.group_size, F = a.n_features;
    for (int g0 = 0; g0 < n; g0 += G) {
        const int n_in = (n - g0) < G ? (n - g0) : G;
        const uint32_t col0 = a.levels[idx[g0]].out_col;
        if (col0 % 4 || (n_in * F) % 4) return false;
        for (int i = 1; i < n_in; ++i)
            if (a.levels[idx[g0 + i]].out_col != col0 + static_cast<uint32_t>(i * F)) return false;
    }
    return true;
}

// Developer tuning knobs (never required; defaults are the measured best).  Each knob is read from the environment
// ONCE per process, at its first use, so a later setenv() cannot silently change the launch geometry of a running job.
int read_env(const char* name, int dflt, int lo, int hi) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    const long x = strtol(v, nullptr, 0);
    return (x < lo || x > hi) ? dflt : static_cast<int>(x);
}
#define DE_KNOB(name, dflt, lo, hi) ([]() -> int { static const int v = read_env(name, dflt, lo, hi); return v; }())

// Ring geometry: ~64 points per stage (a whole number of CTA passes), 4 stages, long-lived CTAs.
void plan_stages(const CallArgs& a, de::EncodeParams& P) {
    const int p_step = de::kThreads / (2 * P.group_size);
    int tp = DE_KNOB("DE_BWD_STAGE_POINTS", 64, 8, 1024);
    tp = ((tp + p_step - 1) / p_step) * p_step;
    P.stage_points = tp;
    P.n_stages = DE_KNOB("DE_BWD_STAGES", 4, 2, 8);
    while (P.n_stages > 2 && de::staged_smem_bytes(P, a.n_features) > 96 * 1024) --P.n_stages;
    long long per_cta = (a.n_points + 148ll * 8 - 1) / (148ll * 8);      // >= 8 CTAs per SM and group
    long long stages = (per_cta + tp - 1) / tp;
    const int max_stages = DE_KNOB("DE_BWD_CTA_STAGES", 32, 1, 4096);
    if (stages < 1) stages = 1;
    if (stages > max_stages) stages = max_stages;
    P.points_per_cta = static_cast<int>(stages) * tp;
    P.n_tiles = static_cast<int>((a.n_points + P.points_per_cta - 1) / P.points_per_cta);
}

constexpr unsigned long long kSmemTableBudget = 200ull << 10;   // of the 227 KB a CTA may own
constexpr long long kSmemMinPoints = 1ll << 16;                 // below this the table staging does not amortise

unsigned long long live_cells(const de_level_desc& L) {
    unsigned long long live = 1;
    for (int d = 0; d < L.n_dims; ++d) live *= (unsigned long long)L.resolution + 1ull;
    return live;
}

// ---- gradient replicas for contended coarse levels (backward) --------------------------------------
// A level whose live entries are few receives thousands of reductions per entry; the L2 atomic units
// serialise same-address updates (profiles/r01c_level_sweep_c3.json: R=16 costs 6x a fine level on
// clustered points).  Such a level gets K copies of its live entries -- CTA tile t adds into copy t % K --
// which are folded into the real table by one small kernel afterwards.
constexpr unsigned long long kRepTargetBytes = 8ull << 20;   // replicas of one level stay within this much L2
constexpr unsigned long long kRepGroupBytes = 16ull << 20;   // ... and the copies of one level group within this much
constexpr unsigned kRepMax = 64;
constexpr unsigned long long kRepMinUpdatesPerEntry = 256;

// ---- nearest-cell 1-D levels served from shared memory (grid4d_small1d_kernels.cu) ---------------------------------
// prime[0] = 1 (hash_encoding.py:64) makes a 1-D level's index the cell number: coordinates in [0, 1] touch rows 0 .. R.
bool small1d_level(const de_level_desc& L, bool linear, int n_features) {
    return !linear && L.n_dims == 1 && (n_features == 1 || n_features == 2 || n_features == 4 || n_features == 8);
}
constexpr unsigned kSmall1dBudget = 40u << 10;      // shared memory per CTA spent on live rows / accumulators
bool plan_small1d(const de_level_desc* levels, int begin, int end, int n_features, int elem_bytes, bool backward,
                  de::Small1dPlan& S) {
    const int n = end - begin;
    if (n > de::kSmall1dMaxLevels) return false;
    S = de::Small1dPlan{};
    const int slots_per_warp = 32 / (n <= 1 ? 1 : (n <= 2 ? 2 : (n <= 4 ? 4 : (n <= 8 ? 8 : 16))));
    unsigned used = 0;
    bool any = false;
    for (int i = 0; i < n; ++i) {                    // coarse levels first: they are the contended ones
        const de_level_desc& L = levels[begin + i];
        unsigned long long live = (unsigned long long)L.resolution + 1ull;
        if (live > L.table_size) live = L.table_size;
        unsigned copies = 1;
        if (backward) { copies = slots_per_warp > 4 ? 4 : slots_per_warp; }
        unsigned long long bytes = (live * n_features * elem_bytes * copies + 15ull) & ~15ull;
        while (copies > 1 && used + bytes > kSmall1dBudget) { copies >>= 1; bytes = (live * n_features * elem_bytes * copies + 15ull) & ~15ull; }
        S.off[i] = used;
        S.copies[i] = copies;
        if (used + bytes > kSmall1dBudget) { S.live[i] = 0; continue; }
        S.live[i] = static_cast<uint32_t>(live);
        used += static_cast<unsigned>(bytes);
        any = true;
    }
    S.total_bytes = used;
    return any;
}

struct RepPlan {
    unsigned count;      // replicas including the table itself (1 = none)
    unsigned live;       // entries per replica
    unsigned stride;     // floats between replicas
};

RepPlan plan_replicas(const de_level_desc& L, int n_features, int64_t n_points, bool linear) {
    RepPlan r{1, 0, 0};
    if (small1d_level(L, linear, n_features)) return r;   // accumulated in shared memory instead (plan_small1d); a 1-D level
                                                          // too fine for that has no contended rows either
    unsigned long long live = L.dense ? live_cells(L) : L.table_size;
    if (!L.dense) {
        // a hashed level touches min(T, (R+1)^D) entries scattered over the whole table -- except in 1-D,
        // where prime[0] = 1 (hash_encoding.py:64) makes the index the cell number itself: entries 0..R
        live = L.table_size;
        if (L.n_dims == 1 && (unsigned long long)L.resolution + 1ull < live) live = (unsigned long long)L.resolution + 1ull;
    }
    const unsigned long long bytes = live * n_features * 4ull;
    const unsigned long long corners = linear ? (1ull << L.n_dims) : 1ull;
    if (bytes == 0 || bytes * 2 > kRepTargetBytes) return r;
    if ((unsigned long long)n_points * corners < kRepMinUpdatesPerEntry * live) return r;
    unsigned long long k = kRepTargetBytes / bytes;
    if (k > kRepMax) k = kRepMax;
    if (k < 2) return r;
    r.count = static_cast<unsigned>(k);
    r.live = static_cast<unsigned>(live);
    r.stride = static_cast<unsigned>((live * n_features + 63ull) & ~63ull);
    return r;
}

size_t replica_workspace_bytes(const de_level_desc* levels, int n_levels, int n_features, int64_t n_points, bool linear) {
    size_t total = 0;
    for (int i = 0; i < n_levels; ++i) {
        const RepPlan r = plan_replicas(levels[i], n_features, n_points, linear);
        if (r.count > 1) total += static_cast<size_t>(r.count - 1) * r.stride * 4;
    }
    return total;
}

// Levels per CTA group of the pair kernels: as many as keep the group's (gradient) tables L2 resident, up to the 16
// that make a point's slice of the feature row a full 128 B line (F = 2).
unsigned long long level_bytes(const de_level_desc& L, int n_features, bool backward, int table_dtype) {
    const unsigned long long live = L.dense ? live_cells(L) : L.table_size;      // a dense level only touches (R+1)^D cells
    const unsigned long long eb = (unsigned long long)n_features * (backward ? 4u : (table_dtype == DE_F16 ? 2u : 4u));
    return live * eb;
}
int pick_group_size(bool pair, int n_features, int levels_per_group, int n, unsigned long long max_level_bytes,
                    bool nearest = false) {
    int G = levels_per_group;
    if (G == 0) {
        if (pair) {
            G = 16;
            while (G > 1 && (unsigned long long)G * max_level_bytes > kL2BudgetBytes) G >>= 1;
        } else {
            // the G lanes of a point cover one 32 B sector of its row; two sectors in nearest-cell mode, where a
            // lane issues ONE gather per point and the coordinate re-read per group is a visible share
            // (c1r, run ak: forward 0.675 -> 0.595 ms, backward 0.966 -> 0.811 ms; 16 = same, 2 = far slower)
            G = (nearest ? 16 : 8) / n_features;
            if (nearest && G > 1 && (unsigned long long)G * max_level_bytes > kL2BudgetBytes) G >>= 1;   // as before
            if (G < 1) G = 1;
        }
    }
    G = floor_pow2(G > n ? n : G);
    if (G > de::kMaxGroup) G = de::kMaxGroup;
    return G;
}

// Backward of one launch (levels [begin, end), equal D): when a group's slice of a gradient row is at most 64 B the
// rows are first re-laid group-major in the workspace (regroup_kernel, misc.cu), so that each walk over the points
// streams contiguous bytes instead of one 128 B HBM line per 16..64 B slice.
struct RegroupPlan {
    bool on;
    int G, W;           // levels per group, floats per slice
    size_t bytes;       // workspace the regrouped rows take
};
constexpr long long kRegroupMinPoints = 1ll << 20;
RegroupPlan plan_regroup(const de_level_desc* levels, int begin, int end, int n_features, int64_t n_points, bool linear,
                         uint32_t flags, int levels_per_group) {
    RegroupPlan r{false, 0, 0, 0};
    const int n = end - begin;
    if (!de::pair_layout_supported(linear, n_features) || (flags & (DE_FLAG_STAGED_BWD | DE_FLAG_NO_REGROUP))) return r;
    if (n_points < kRegroupMinPoints) return r;
    unsigned long long max_level_bytes = 1;
    for (int i = begin; i < end; ++i) {
        const unsigned long long b = level_bytes(levels[i], n_features, true, DE_F32);
        if (b > max_level_bytes) max_level_bytes = b;
        if (levels[i].out_col != levels[begin].out_col + static_cast<uint32_t>((i - begin) * n_features)) return r;   // columns must be consecutive
    }
    r.G = pick_group_size(true, n_features, levels_per_group, n, max_level_bytes);
    r.W = r.G * n_features;
    const int n_cols = n * n_features;
    if (r.W % 4 || r.W * 4 > 64 || n % r.G || n / r.G < 4) return r;
    if ((n_cols + 4) * 32 * 4 > 48 * 1024 || levels[begin].out_col % 4) return r;
    r.bytes = (static_cast<size_t>(n_cols) * static_cast<size_t>(n_points) * 4 + 255) & ~static_cast<size_t>(255);
    r.on = true;
    return r;
}
size_t regroup_workspace_bytes(const de_level_desc* levels, int n_levels, int n_features, int64_t n_points, bool linear,
                               uint32_t flags, int levels_per_group) {
    if (flags & DE_FLAG_GRAD_GROUPED) return 0;         // the producer wrote the group-major layout itself
    size_t total = 0;
    int begin = 0;
    while (begin < n_levels) {
        int end = begin + 1;
        while (end < n_levels && levels[end].n_dims == levels[begin].n_dims && end - begin < de::kMaxLaunchLevels) ++end;
        const RegroupPlan r = plan_regroup(levels, begin, end, n_features, n_points, linear, flags, levels_per_group);
        if (r.on) total += r.bytes;
        begin = end;
    }
    return total;
}

// Forward only: picks the dense levels of [begin, end) whose live cells fit shared memory, packs them into
// shared-memory groups and launches grid4d_forward_smem_kernel for them.  taken[i] marks the levels served.
int run_smem_forward(const CallArgs& a, const de::EncodeParams& base, int begin, int end, bool* taken,
                     cudaStream_t stream) {
    const int n = end - begin;
    for (int i = 0; i < n; ++i) taken[i] = false;
    const bool linear = (a.flags & DE_INTERP_MASK) == DE_INTERP_LINEAR;
    const bool half_tables = a.table_dtype == DE_F16;
    if (!linear || !(a.flags & DE_FLAG_SMEM_COARSE) || a.n_points < kSmemMinPoints) return DE_OK;
    if (!de::smem_layout_supported(half_tables, a.n_features)) return DE_OK;
    const unsigned long long eb = (unsigned long long)a.n_features * (half_tables ? 2u : 4u);

    de::EncodeParams P = base;
    de::SmemPlan S{};
    int n_lv = 0;
    bool pow2_all = true;
    unsigned long long used = 0;
    for (int i = 0; i < n; ++i) {
        const de_level_desc& L = a.levels[begin + i];
        if (!L.dense) continue;
        const unsigned long long bytes = (live_cells(L) * eb + 15ull) & ~15ull;
        if (bytes > kSmemTableBudget) continue;
        if (reinterpret_cast<uintptr_t>(L.table) % 4) continue;
        int g = S.n_sgroups - 1;
        // a group holds consecutive picks of ONE grid (same coordinate columns), <= 4 levels, <= budget
        bool join = false;
        if (g >= 0 && S.count[g] < de::kSmemGroupLevels && used + bytes <= kSmemTableBudget) {
            // the group's first level sits at P.levels[S.first[g]]; compare its dims with L's
            const de::LevelParams& f = P.levels[S.first[g]];
            join = f.mask_id == L.mask_id;
            for (int d = 0; d < L.n_dims && join; ++d) join = f.dim[d] == L.dim[d];
        }
        if (!join) {
            if (S.n_sgroups == de::kMaxSmemGroups) break;
            g = S.n_sgroups++;
            S.first[g] = n_lv;
            S.count[g] = 0;
            used = 0;
        }
        fill_level(L, P.levels[n_lv]);
        S.off[g][S.count[g]] = static_cast<uint32_t>(used);
        S.live[g][S.count[g]] = static_cast<uint32_t>(live_cells(L));
        ++S.count[g];
        used += bytes;
        if (used > S.smem_bytes) S.smem_bytes = static_cast<unsigned>(used);
        pow2_all = pow2_all && is_pow2(L.table_size);
        taken[i] = true;
        ++n_lv;
    }
    if (n_lv == 0) return DE_OK;

    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    int cta = 0;
    for (int g = 0; g < S.n_sgroups; ++g) {
        int gs = 1;
        while (gs < S.count[g]) gs <<= 1;
        S.gs[g] = gs;
        int c = static_cast<int>(static_cast<long long>(n_sms) * S.count[g] / n_lv);   // one CTA per SM in total
        if (c < 1) c = 1;
        S.cta_begin[g] = cta;
        S.cta_count[g] = c;
        cta += c;
    }
    S.n_ctas = cta;
    P.n_levels = n_lv;
    cudaError_t e = de::launch_forward_smem(P, S, a.levels[begin].n_dims, half_tables, a.n_features, pow2_all, stream);
    if (e != cudaSuccess) return cuda_fail(e, "shared-memory forward launch");
    return DE_OK;
}

// Cuts the level list into launches (runs of equal D, <= 64 levels) and starts them in order.
int run(const CallArgs& a, Op op, de::EncodeParams& base, cudaStream_t stream) {
    if (a.n_points == 0) return DE_OK;
    const bool linear = (a.flags & DE_INTERP_MASK) == DE_INTERP_LINEAR;
    size_t ws_off = 0;
    size_t regroup_off = 0;         // regrouped gradient rows live behind the replicas
    if (op == Op::Backward && a.workspace) {
        const size_t used = replica_workspace_bytes(a.levels, a.n_levels, a.n_features, a.n_points, linear);
        // replicas are granted front to back while they fit (below), so every granted range lies inside
        // [0, min(used, workspace_bytes)): zero exactly that, also when the caller's scratch is smaller than asked for
        const size_t zeroed = used < a.workspace_bytes ? used : a.workspace_bytes;
        if (zeroed > 0) {
            cudaError_t e = cudaMemsetAsync(a.workspace, 0, zeroed, stream);
            if (e != cudaSuccess) return cuda_fail(e, "workspace memset");
        }
        regroup_off = (used + 255) & ~static_cast<size_t>(255);
    }
    int begin = 0;
    while (begin < a.n_levels) {
        const int D = a.levels[begin].n_dims;
        int end = begin + 1;
        while (end < a.n_levels && a.levels[end].n_dims == D && end - begin < de::kMaxLaunchLevels) ++end;
        const bool pair = op != Op::Indices && de::pair_layout_supported(linear, a.n_features);

        bool taken[de::kMaxLaunchLevels] = {};
        if (op == Op::Forward && pair)
            if (int rc = run_smem_forward(a, base, begin, end, taken, stream)) return rc;

        de::EncodeParams P = base;
        P.l2_hints = (a.flags & DE_FLAG_NO_L2_HINTS) ? 0 : ((a.flags & DE_FLAG_COHERENT_GATHER) ? 2 : 1);
        int idx[de::kMaxLaunchLevels];
        int n = 0;
        for (int i = begin; i < end; ++i)
            if (!taken[i - begin]) idx[n++] = i;
        if (n == 0) { begin = end; continue; }
        P.n_levels = n;
        bool pow2_all = true;
        unsigned long long max_level_bytes = 1;
        de::ReplicaFoldParams fold{};
        for (int i = 0; i < n; ++i) {
            const de_level_desc& L = a.levels[idx[i]];
            fill_level(L, P.levels[i]);
            pow2_all = pow2_all && is_pow2(L.table_size);
            const unsigned long long lb = level_bytes(L, a.n_features, op == Op::Backward, a.table_dtype);
            if (lb > max_level_bytes) max_level_bytes = lb;
        }
        const int G = pick_group_size(pair, a.n_features, a.levels_per_group, n, max_level_bytes, !linear);
        if (op == Op::Backward && a.workspace) {
            // Replicas are extra L2 footprint of the group being walked: grant them coarse level first, while one
            // group's copies stay within kRepGroupBytes.  (Without the cap, 2^26 points make every T = 2^19 level
            // "contended" by the updates-per-entry rule, each gets one 4 MiB copy, the group's 64 MiB becomes 116 MiB
            // and the backward runs 2.7x slower -- bench c4, run as.)
            unsigned long long group_extra = 0;
            for (int i = 0; i < n; ++i) {
                if (i % G == 0) group_extra = 0;
                const de_level_desc& L = a.levels[idx[i]];
                const RepPlan r = plan_replicas(L, a.n_features, a.n_points, linear);
                if (r.count <= 1) continue;
                const size_t need = static_cast<size_t>(r.count - 1) * r.stride * 4;
                if (group_extra + need > kRepGroupBytes || ws_off + need > a.workspace_bytes) continue;
                de::LevelParams& lp = P.levels[i];
                lp.rep_base = reinterpret_cast<float*>(static_cast<char*>(a.workspace) + ws_off);
                lp.rep_count = r.count;
                lp.rep_stride = r.stride;
                lp.rep_live = r.live;
                fold.dst[fold.n] = L.grad_table;
                fold.rep[fold.n] = lp.rep_base;
                fold.n_floats[fold.n] = r.live * static_cast<unsigned>(a.n_features);
                fold.count[fold.n] = r.count - 1;
                fold.stride[fold.n] = r.stride;
                ++fold.n;
                ws_off += need;
                group_extra += need;
            }
        }
        if (op == Op::Backward && (a.flags & DE_FLAG_GRAD_GROUPED)) {
            // the caller (the head's backward kernel) already wrote the rows group-major: same addressing, no pass
            const RegroupPlan rg = plan_regroup(a.levels, begin, end, a.n_features, a.n_points, linear,
                                                a.flags & ~DE_FLAG_NO_REGROUP, a.levels_per_group);
            if (!(begin == 0 && end == a.n_levels && pair && rg.on && rg.G == G && base.out_stride == rg.W))
                return fail(DE_EINVAL, "DE_FLAG_GRAD_GROUPED: grad_out layout (W=%lld) does not match this call's level groups "
                            "(de_grid4d_backward_group_width)", (long long)base.out_stride);
            P.vec_rows = 1;
            for (int i = 0; i < n; ++i)
                P.levels[i].out_col = static_cast<unsigned long long>(i / G) * a.n_points * rg.W + (i % G) * a.n_features;
        } else if (op == Op::Backward && a.workspace && pair) {
            const RegroupPlan rg = plan_regroup(a.levels, begin, end, a.n_features, a.n_points, linear, a.flags, a.levels_per_group);
            if (rg.on && rg.G == G && regroup_off + rg.bytes <= a.workspace_bytes &&
                reinterpret_cast<uintptr_t>(base.grad_out) % 16 == 0 && base.out_stride % 4 == 0) {
                float* dst = reinterpret_cast<float*>(static_cast<char*>(a.workspace) + regroup_off);
                int dev = 0, n_sms = 148;
                if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
                cudaError_t e = de::launch_regroup(base.grad_out, base.out_stride, static_cast<int>(a.levels[begin].out_col),
                                                   n * a.n_features, rg.W, dst, a.n_points, n_sms, stream);
                if (e != cudaSuccess) return cuda_fail(e, "regroup launch");
                // the backward kernels address grad_out + p * out_stride + out_col: plane g = level / G starts at
                // g * n_points * W floats, rows are W floats apart
                P.grad_out = dst;
                P.out_stride = rg.W;
                P.vec_rows = 1;
                for (int i = 0; i < n; ++i)
                    P.levels[i].out_col = static_cast<unsigned long long>(i / G) * a.n_points * rg.W + (i % G) * a.n_features;
                regroup_off += rg.bytes;
            }
        }
        if ((op == Op::Forward || op == Op::Backward) && n == end - begin && small1d_level(a.levels[begin], linear, a.n_features) &&
            !(a.flags & DE_FLAG_NO_SMALL1D)) {
            de::Small1dPlan S;
            const int eb = (op == Op::Forward && a.table_dtype == DE_F16) ? 2 : 4;
            if (plan_small1d(a.levels, begin, end, a.n_features, eb, op == Op::Backward, S)) {
                int G1 = 1;
                while (G1 < n) G1 <<= 1;
                P.group_size = G1;
                P.n_groups = 1;
                const long long p_step = de::kThreads / G1;
                long long ppc = (a.n_points + 148 * 4 - 1) / (148 * 4);      // a few long-lived CTAs per SM: one flush each
                if (ppc < p_step * 8) ppc = p_step * 8;
                ppc = ((ppc + p_step - 1) / p_step) * p_step;
                if (ppc > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "n_points=%lld too large for one launch", (long long)a.n_points);
                P.points_per_cta = static_cast<int>(ppc);
                const long long ctas = (a.n_points + ppc - 1) / ppc;
                P.n_tiles = static_cast<int>(ctas);
                cudaError_t e1 = op == Op::Forward
                    ? de::launch_small1d_forward(P, S, a.table_dtype == DE_F16, a.n_features, static_cast<int>(ctas), stream)
                    : de::launch_small1d_backward(P, S, a.n_features, static_cast<int>(ctas), stream);
                if (e1 != cudaSuccess) return cuda_fail(e1, "shared-memory 1-D launch");
                begin = end;
                continue;
            }
        }
        P.group_size = G;
        P.n_groups = (n + G - 1) / G;
        const int lanes_per_point = pair ? 2 * G : G;
        // loop trips per CTA (run ao / ap): the forward kernels want long-lived CTAs (set-up amortised, fewer block
        // launches), the backward kernels do not (shorter tiles spread the reductions of a group more evenly)
        const bool fwd_like = op != Op::Backward;
        const int iters = pair ? (fwd_like ? DE_KNOB("DE_PAIR_TILE_ITERS_FWD", 32, 1, 4096) : DE_KNOB("DE_PAIR_TILE_ITERS_BWD", 8, 1, 4096))
                               : (fwd_like ? DE_KNOB("DE_TILE_ITERS_FWD", 16, 1, 4096) : DE_KNOB("DE_TILE_ITERS_BWD", 8, 1, 4096));
        int trips = iters;
        {
            // small batches: never fewer CTAs than the short tiles of r01h gave (8 / 4 trips) until there are a few
            // waves of them (148 SMs x 8 resident CTAs x 4)
            const int floor_trips = pair ? 8 : 4;
            const long long per_pass = de::kThreads / lanes_per_point;
            while (trips > floor_trips && ((a.n_points + per_pass * trips - 1) / (per_pass * trips)) * P.n_groups < 4736) trips >>= 1;
        }
        P.points_per_cta = (de::kThreads / lanes_per_point) * trips;
        const long long tiles = (a.n_points + P.points_per_cta - 1) / P.points_per_cta;
        if (tiles * P.n_groups >= (1ll << 31)) return fail(DE_EUNSUPPORTED, "n_points=%lld too large for one launch", (long long)a.n_points);
        P.n_tiles = static_cast<int>(tiles);

        cudaError_t e = cudaSuccess;
        switch (op) {
            case Op::Forward:
                e = pair ? de::launch_forward_pair(P, D, a.table_dtype == DE_F16, a.n_features, pow2_all, stream)
                         : de::launch_forward(P, D, linear, a.table_dtype == DE_F16, a.n_features, stream);
                break;
            case Op::Backward:
                if (pair && staging_ok(a, P, idx, n)) {
                    plan_stages(a, P);
                    e = de::launch_backward_staged(P, D, a.n_features, pow2_all, stream);
                } else {
                    e = pair ? de::launch_backward_pair(P, D, a.n_features, pow2_all, stream)
                             : de::launch_backward(P, D, linear, a.n_features, stream);
                }
                break;
            case Op::Indices: {
                // indices are laid out per call, so offset the level index by `begin`
                const int C = linear ? (1 << D) : 1;
                P.indices = base.indices + static_cast<long long>(begin) * C * a.n_points;
                e = de::launch_indices(P, D, linear, stream);
                break;
            }
        }
        if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
        if (fold.n) {
            e = de::launch_replica_fold(fold, stream);
            if (e != cudaSuccess) return cuda_fail(e, "replica fold launch");
        }
        begin = end;
    }
    return DE_OK;
}

}  // namespace

extern "C" {

DE_API int de_abi_version(void) { return DE_ABI_VERSION; }

DE_API const char* de_last_error(void) { return g_err; }

DE_API int de_corners(uint32_t flags, int n_dims) {
    return ((flags & DE_INTERP_MASK) == DE_INTERP_LINEAR) ? (1 << n_dims) : 1;
}

DE_API int de_grid4d_hash_indices(const de_level_desc* levels_host, int n_levels, const float* coords,
                           int64_t n_points, int coord_stride, uint32_t flags, int64_t* indices,
                           void* stream) {
    CallArgs a{levels_host, n_levels, coords, n_points, coord_stride, DE_F32, 1, flags, 0, false, false};
    if (int rc = validate(a)) return rc;
    for (int i = 1; i < n_levels; ++i)
        if (levels_host[i].n_dims != levels_host[0].n_dims) return fail(DE_EINVAL, "hash_indices: all levels must share n_dims");
    if (n_points > 0 && !indices) return fail(DE_EINVAL, "indices is NULL");
    de::EncodeParams P{};
    P.coords = coords;
    P.indices = reinterpret_cast<long long*>(indices);
    P.n_points = n_points;
    P.coord_stride = coord_stride;
    P.n_features = 1;
    P.vec_coords = (coord_stride == 4 && reinterpret_cast<uintptr_t>(coords) % 16 == 0) ? 1 : 0;
    P.out_scale = 1.0f;
    return run(a, Op::Indices, P, static_cast<cudaStream_t>(stream));
}

DE_API int de_grid4d_forward(const de_level_desc* levels_host, int n_levels, const float* coords,
                      int64_t n_points, int coord_stride, int table_dtype, int n_features,
                      float out_scale, const uint8_t* mask_a, const uint8_t* mask_b, float* out,
                      int64_t out_stride, uint32_t flags, int levels_per_group, void* stream) {
    CallArgs a{levels_host, n_levels, coords, n_points, coord_stride, table_dtype, n_features, flags, levels_per_group, true, false};
    if (int rc = validate(a)) return rc;
    if (n_points > 0 && !out) return fail(DE_EINVAL, "out is NULL");
    for (int i = 0; i < n_levels; ++i)
        if ((int64_t)levels_host[i].out_col + n_features > out_stride)
            return fail(DE_EINVAL, "level %d: out_col %u + F %d exceeds out_stride %lld", i, levels_host[i].out_col, n_features, (long long)out_stride);
    de::EncodeParams P{};
    P.coords = coords;
    P.mask_a = mask_a;
    P.mask_b = mask_b;
    P.out = out;
    P.n_points = n_points;
    P.out_stride = out_stride;
    P.coord_stride = coord_stride;
    P.n_features = n_features;
    P.vec_coords = (coord_stride == 4 && reinterpret_cast<uintptr_t>(coords) % 16 == 0) ? 1 : 0;
    P.vec_rows = rows_vectorisable(out, out_stride, levels_host, n_levels, n_features) ? 1 : 0;
    P.out_scale = out_scale;
    return run(a, Op::Forward, P, static_cast<cudaStream_t>(stream));
}

DE_API int de_grid4d_backward(const de_level_desc* levels_host, int n_levels, const float* coords,
                       int64_t n_points, int coord_stride, int n_features, const uint8_t* mask_a,
                       const uint8_t* mask_b, const float* grad_out, int64_t grad_stride,
                       uint32_t flags, int levels_per_group, void* workspace, size_t workspace_bytes,
                       void* stream) {
    CallArgs a{levels_host, n_levels, coords, n_points, coord_stride, DE_F32, n_features, flags, levels_per_group, false, true};
    a.workspace = workspace_bytes ? workspace : nullptr;
    a.workspace_bytes = workspace_bytes;
    if (a.workspace && reinterpret_cast<uintptr_t>(a.workspace) % 16) return fail(DE_EALIGN, "workspace not 16-byte aligned");
    if (int rc = validate(a)) return rc;
    if (n_points > 0 && !grad_out) return fail(DE_EINVAL, "grad_out is NULL");
    for (int i = 0; i < n_levels && !(flags & DE_FLAG_GRAD_GROUPED); ++i)
        if ((int64_t)levels_host[i].out_col + n_features > grad_stride)
            return fail(DE_EINVAL, "level %d: out_col %u + F %d exceeds grad_stride %lld", i, levels_host[i].out_col, n_features, (long long)grad_stride);
    if ((flags & DE_FLAG_GRAD_GROUPED) && reinterpret_cast<uintptr_t>(grad_out) % 16) return fail(DE_EALIGN, "grouped grad_out not 16-byte aligned");
    de::EncodeParams P{};
    P.coords = coords;
    P.mask_a = mask_a;
    P.mask_b = mask_b;
    P.grad_out = grad_out;
    P.n_points = n_points;
    P.out_stride = grad_stride;
    P.coord_stride = coord_stride;
    P.n_features = n_features;
    P.vec_coords = (coord_stride == 4 && reinterpret_cast<uintptr_t>(coords) % 16 == 0) ? 1 : 0;
    P.vec_rows = rows_vectorisable(grad_out, grad_stride, levels_host, n_levels, n_features) ? 1 : 0;
    P.out_scale = 1.0f;
    return run(a, Op::Backward, P, static_cast<cudaStream_t>(stream));
}

DE_API size_t de_grid4d_sort_workspace_bytes(int n_dims, int bits) {
    if (n_dims < 1 || n_dims > 4 || bits < 1 || bits * n_dims > 16) return 0;
    return static_cast<size_t>(4) << (bits * n_dims);
}

DE_API int de_grid4d_sort_points(const float* coords, int64_t n_points, int coord_stride, const uint8_t* dims_host, int n_dims,
                          int bits, int32_t* perm, void* workspace, size_t workspace_bytes, void* stream) {
    if (n_points < 0 || n_points > 0x7fffffffll) return fail(DE_EINVAL, "sort_points: n_points=%lld", (long long)n_points);
    if (n_points == 0) return DE_OK;
    if (!coords || !dims_host || !perm || !workspace) return fail(DE_EINVAL, "sort_points: NULL pointer");
    const size_t need = de_grid4d_sort_workspace_bytes(n_dims, bits);
    if (!need) return fail(DE_EUNSUPPORTED, "sort_points: n_dims=%d bits=%d (bits * n_dims must be in 1..16)", n_dims, bits);
    if (workspace_bytes < need) return fail(DE_EINVAL, "sort_points: workspace of %zu bytes, %zu needed", workspace_bytes, need);
    de::SortParams S{};
    S.coords = coords; S.n = n_points; S.coord_stride = coord_stride; S.n_dims = n_dims; S.bits = bits;
    for (int d = 0; d < n_dims; ++d) {
        if (dims_host[d] >= coord_stride) return fail(DE_EINVAL, "sort_points: dim[%d]=%d >= coord_stride=%d", d, dims_host[d], coord_stride);
        S.dim[d] = dims_host[d];
    }
    S.counts = static_cast<unsigned int*>(workspace);
    S.perm = perm;
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e = de::launch_sort_points(S, n_sms, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "sort_points launch");
}

// ---- de_grid4d_sort_points without any allocation on the caller's side ---------------------------------------------------
// The counting sort's bucket counters live in a file-scope device array instead of a caller-provided workspace: 2^16 buckets
// (the entry point's own limit, bits * n_dims <= 16) x 4 bytes = 256 KiB, held by every GPU that loads the library until the
// process exits.  `g_sort_counters` is, in host code, only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current, and is what the kernels receive.  The one piece of
// mutable global state in this file (the rest of the surface is re-entrant): calls share the array, see the header.
namespace {

constexpr int kNoAllocSortMaxBuckets = DE_SORT_NOALLOC_MAX_BUCKETS;
static __device__ __align__(256) unsigned int g_sort_counters[kNoAllocSortMaxBuckets];

constexpr int kNoAllocMaxDevices = 64;
unsigned int* g_sort_counters_addr[kNoAllocMaxDevices];           // [device ordinal] -> device address of g_sort_counters there

int sort_counters_for_current_device(unsigned int** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "sort_points_noalloc: cudaGetDevice");
    if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(DE_EUNSUPPORTED, "sort_points_noalloc: device ordinal %d >= %d", dev, kNoAllocMaxDevices);
    unsigned int* p = __atomic_load_n(&g_sort_counters_addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        // set-up, once per device (two threads racing here obtain the same address)
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_sort_counters);
        if (e != cudaSuccess) return cuda_fail(e, "sort_points_noalloc: cudaGetSymbolAddress");
        if (!p) return fail(DE_EINVAL, "sort_points_noalloc: cudaGetSymbolAddress returned a null address");
        __atomic_store_n(&g_sort_counters_addr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return DE_OK;
}

}  // namespace

DE_API int de_grid4d_sort_points_noalloc(const float* coords, int64_t n_points, int coord_stride, const uint8_t* dims_host, int n_dims,
                                  int bits, int32_t* perm, void* stream) {
    if (n_points < 0 || n_points > 0x7fffffffll) return fail(DE_EINVAL, "sort_points_noalloc: n_points=%lld", (long long)n_points);
    if (n_dims < 1 || n_dims > 4 || bits < 1 || bits * n_dims > 16 || (1ll << (bits * n_dims)) > kNoAllocSortMaxBuckets)
        return fail(DE_EUNSUPPORTED, "sort_points_noalloc: n_dims=%d bits=%d needs 2^%d buckets; the built-in scratch holds %d (bits * n_dims must be in 1..16)",
                    n_dims, bits, bits * n_dims, kNoAllocSortMaxBuckets);
    unsigned int* counters = nullptr;
    if (int rc = sort_counters_for_current_device(&counters)) return rc;      // also with n_points = 0: a caller's warm-up call
    if (n_points == 0) return DE_OK;
    if (!coords || !dims_host || !perm) return fail(DE_EINVAL, "sort_points_noalloc: NULL pointer");
    de::SortParams S{};
    S.coords = coords; S.n = n_points; S.coord_stride = coord_stride; S.n_dims = n_dims; S.bits = bits;
    for (int d = 0; d < n_dims; ++d) {
        if (dims_host[d] >= coord_stride) return fail(DE_EINVAL, "sort_points_noalloc: dim[%d]=%d >= coord_stride=%d", d, dims_host[d], coord_stride);
        S.dim[d] = dims_host[d];
    }
    S.counts = counters;          // launch_sort_points zeroes the 2^(bits * n_dims) counters it uses (cudaMemsetAsync on `stream`)
    S.perm = perm;
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e = de::launch_sort_points(S, n_sms, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "sort_points_noalloc launch");
}

DE_API int de_grid4d_backward_sorted(const de_level_desc* levels_host, int n_levels, const float* coords, int64_t n_points,
                              int coord_stride, int n_features, const uint8_t* mask_a, const uint8_t* mask_b,
                              const float* grad_out, int64_t grad_stride, uint32_t flags, const int32_t* perm, void* stream) {
    CallArgs a{levels_host, n_levels, coords, n_points, coord_stride, DE_F32, n_features, flags, 0, false, true};
    if (int rc = validate(a)) return rc;
    if (n_points == 0) return DE_OK;
    if (!grad_out || !perm) return fail(DE_EINVAL, "backward_sorted: NULL pointer");
    if ((flags & DE_INTERP_MASK) != DE_INTERP_LINEAR) return fail(DE_EUNSUPPORTED, "backward_sorted: DE_INTERP_LINEAR levels only");
    if (!(n_features == 1 || n_features == 2 || n_features == 4)) return fail(DE_EUNSUPPORTED, "backward_sorted: n_features=%d not in {1,2,4}", n_features);
    if (n_levels > de::kMaxLaunchLevels) return fail(DE_EUNSUPPORTED, "backward_sorted: n_levels=%d > %d", n_levels, de::kMaxLaunchLevels);
    de::EncodeParams P{};
    P.coords = coords; P.mask_a = mask_a; P.mask_b = mask_b; P.grad_out = grad_out; P.n_points = n_points;
    P.out_stride = grad_stride; P.coord_stride = coord_stride; P.n_features = n_features;
    P.vec_coords = (coord_stride == 4 && reinterpret_cast<uintptr_t>(coords) % 16 == 0) ? 1 : 0;
    P.vec_rows = rows_vectorisable(grad_out, grad_stride, levels_host, n_levels, n_features) ? 1 : 0;
    P.out_scale = 1.0f;
    P.n_levels = n_levels;
    bool pow2_all = true;
    for (int i = 0; i < n_levels; ++i) {
        const de_level_desc& L = levels_host[i];
        if (L.n_dims != 3) return fail(DE_EUNSUPPORTED, "backward_sorted: level %d has n_dims=%d (3-D grids only)", i, L.n_dims);
        if ((int64_t)L.out_col + n_features > grad_stride) return fail(DE_EINVAL, "level %d: out_col %u + F %d exceeds grad_stride %lld", i, L.out_col, n_features, (long long)grad_stride);
        fill_level(L, P.levels[i]);
        pow2_all = pow2_all && is_pow2(L.table_size);
    }
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e = de::launch_backward_sorted(P, perm, n_features, pow2_all, n_sms, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "backward_sorted launch");
}

DE_API size_t de_grid4d_backward_workspace_bytes(const de_level_desc* levels_host, int n_levels, int n_features,
                                                 int64_t n_points, uint32_t flags) {
    return de_grid4d_backward_workspace_bytes_ex(levels_host, n_levels, n_features, n_points, flags, 0);
}

DE_API size_t de_grid4d_backward_workspace_bytes_ex(const de_level_desc* levels_host, int n_levels, int n_features,
                                                    int64_t n_points, uint32_t flags, int levels_per_group) {
    if (!levels_host || n_levels <= 0 || n_features < 1 || n_points <= 0) return 0;
    const bool linear = (flags & DE_INTERP_MASK) == DE_INTERP_LINEAR;
    const size_t rep = replica_workspace_bytes(levels_host, n_levels, n_features, n_points, linear);
    const size_t rg = regroup_workspace_bytes(levels_host, n_levels, n_features, n_points, linear, flags, levels_per_group);
    return rg ? ((rep + 255) & ~static_cast<size_t>(255)) + rg : rep;
}

DE_API int de_grid4d_backward_group_width(const de_level_desc* levels_host, int n_levels, int n_features,
                                          int64_t n_points, uint32_t flags, int levels_per_group) {
    if (!levels_host || n_levels <= 0 || n_levels > de::kMaxLaunchLevels || n_features < 1 || n_points <= 0) return 0;
    for (int i = 1; i < n_levels; ++i)
        if (levels_host[i].n_dims != levels_host[0].n_dims) return 0;           // one launch only
    const RegroupPlan rg = plan_regroup(levels_host, 0, n_levels, n_features, n_points, (flags & DE_INTERP_MASK) == DE_INTERP_LINEAR,
                                        flags & ~DE_FLAG_NO_REGROUP, levels_per_group);
    return rg.on && levels_host[0].out_col == 0 ? rg.W : 0;
}

DE_API int de_tables_to_f16(const de_level_desc* levels_host, int n_levels, int n_features,
                     void* const* dst_tables_host, float scale, void* stream) {
    if (!levels_host || !dst_tables_host || n_levels <= 0) return fail(DE_EINVAL, "tables_to_f16: bad arguments");
    if (n_features < 1) return fail(DE_EINVAL, "n_features=%d", n_features);
    int begin = 0;
    while (begin < n_levels) {
        de::CastParams P{};
        const int n = (n_levels - begin) < 64 ? (n_levels - begin) : 64;
        for (int i = 0; i < n; ++i) {
            if (!levels_host[begin + i].table || !dst_tables_host[begin + i]) return fail(DE_EINVAL, "tables_to_f16: level %d NULL pointer", begin + i);
            P.src[i] = static_cast<const float*>(levels_host[begin + i].table);
            P.dst[i] = dst_tables_host[begin + i];
            P.n_elems[i] = static_cast<long long>(levels_host[begin + i].table_size) * n_features;
        }
        P.n_levels = n;
        P.scale = scale;
        cudaError_t e = de::launch_cast_f16(P, static_cast<cudaStream_t>(stream));
        if (e != cudaSuccess) return cuda_fail(e, "tables_to_f16 launch");
        begin += n;
    }
    return DE_OK;
}

DE_API int de_geo2xyzt_norm(const void* geo, int geo_dtype, const void* t, int t_dtype,
                     const double* bbox_min_host, const double* bbox_span_host,
                     const double* t_range_host, float* xyzt, double* ecef, int64_t n_points,
                     void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (n_points == 0) return DE_OK;
    if (!geo || !xyzt || !bbox_min_host || !bbox_span_host) return fail(DE_EINVAL, "geo2xyzt: NULL pointer");
    if (geo_dtype != DE_F64 && geo_dtype != DE_F32) return fail(DE_EUNSUPPORTED, "geo_dtype=%d", geo_dtype);
    if (t && t_dtype != DE_F64 && t_dtype != DE_F32) return fail(DE_EUNSUPPORTED, "t_dtype=%d", t_dtype);
    if (t && !t_range_host) return fail(DE_EINVAL, "t given without t_range");
    if (reinterpret_cast<uintptr_t>(xyzt) % 16) return fail(DE_EALIGN, "xyzt not 16-byte aligned");
    cudaError_t e = de::launch_geo2xyzt(geo, geo_dtype == DE_F64, t, t_dtype == DE_F64, bbox_min_host,
                                        bbox_span_host, t_range_host, xyzt, ecef, n_points,
                                        static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "geo2xyzt launch");
}

DE_API int de_time_from_components(const void* components, int dtype, int64_t n, int64_t stride, int n_cols,
                            int default_month, int default_day, int default_hour, const double* t_range_host,
                            double* timestamps, float* t_norm, void* stream) {
    if (n < 0) return fail(DE_EINVAL, "n < 0");
    if (n == 0) return DE_OK;
    if (!components || (!timestamps && !t_norm)) return fail(DE_EINVAL, "time_from_components: NULL pointer");
    if (dtype != DE_I32 && dtype != DE_F64) return fail(DE_EUNSUPPORTED, "time_from_components: dtype=%d (DE_I32 or DE_F64)", dtype);
    if (n_cols < 1 || n_cols > 6 || stride < n_cols) return fail(DE_EINVAL, "time_from_components: n_cols=%d stride=%lld", n_cols, (long long)stride);
    if (default_month < 1 || default_month > 12 || default_day < 1 || default_day > 31 || default_hour < 0 || default_hour > 23)
        return fail(DE_EINVAL, "time_from_components: defaults out of range");
    if (t_norm && !t_range_host) return fail(DE_EINVAL, "time_from_components: t_norm without t_range");
    const double t_min = t_range_host ? t_range_host[0] : 0.0;
    const double t_span = t_range_host ? t_range_host[1] - t_range_host[0] : 1.0;
    cudaError_t e = de::launch_time_components(components, dtype == DE_F64, n, stride, n_cols, default_month, default_day,
                                               default_hour, t_min, t_span, timestamps, t_norm, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "time_from_components launch");
}

DE_API int de_geo_bbox(const void* geo, int geo_dtype, double* minmax, int64_t n_points, void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (n_points == 0) return DE_OK;
    if (!geo || !minmax) return fail(DE_EINVAL, "geo_bbox: NULL pointer");
    if (geo_dtype != DE_F64 && geo_dtype != DE_F32) return fail(DE_EUNSUPPORTED, "geo_dtype=%d", geo_dtype);
    cudaError_t e = de::launch_geo_bbox(geo, geo_dtype == DE_F64, minmax, n_points, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "geo_bbox launch");
}

DE_API int de_mlp_head_supported(int in_dim, int hidden, int out_dim) {
    return de::mlp_head_supported(in_dim, hidden, out_dim) ? 1 : 0;
}

DE_API int de_mlp_head_forward(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                        const void* w1_bf16, const float* b1, const float* ln_gamma, const float* ln_beta,
                        float ln_eps, int layernorm, int activation, const void* w2_bf16, const float* b2,
                        int hidden, int out_dim, float* y, int64_t y_stride, void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (n_points == 0) return DE_OK;
    if (!x || !w1_bf16 || !w2_bf16 || !y) return fail(DE_EINVAL, "mlp_head: NULL pointer");
    if (!de::mlp_head_supported(in_dim, hidden, out_dim))
        return fail(DE_EUNSUPPORTED, "mlp_head: in_dim=%d hidden=%d out_dim=%d (built for in_dim in {32,48,64,96,128}, hidden = out = 64)", in_dim, hidden, out_dim);
    if (activation < 0 || activation > 2) return fail(DE_EINVAL, "mlp_head: activation=%d", activation);
    if (x_stride < in_dim || y_stride < out_dim) return fail(DE_EINVAL, "mlp_head: row stride shorter than the row");
    if (reinterpret_cast<uintptr_t>(x) % 16 || x_stride % 4) return fail(DE_EALIGN, "mlp_head: x rows not 16-byte aligned");
    if (reinterpret_cast<uintptr_t>(y) % 16 || y_stride % 4) return fail(DE_EALIGN, "mlp_head: y rows not 16-byte aligned");
    if (reinterpret_cast<uintptr_t>(w1_bf16) % 16 || reinterpret_cast<uintptr_t>(w2_bf16) % 16) return fail(DE_EALIGN, "mlp_head: weights not 16-byte aligned");
    de::MlpHeadParams P{};
    P.x = x; P.n = n_points; P.x_stride = x_stride;
    P.w1 = w1_bf16; P.b1 = b1; P.ln_gamma = ln_gamma; P.ln_beta = ln_beta; P.ln_eps = ln_eps;
    P.layernorm = layernorm ? 1 : 0; P.activation = activation;
    P.w2 = w2_bf16; P.b2 = b2; P.y = y; P.y_stride = y_stride;
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e = de::launch_mlp_head(P, in_dim, n_sms, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "mlp_head launch");
}

DE_API int de_mlp_head_backward(const float* x, const float* grad_y, int64_t n_points, int in_dim, int64_t x_stride,
                         int64_t gy_stride, const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                         const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps, int layernorm,
                         int activation, int hidden, int out_dim, float* grad_x, int64_t gx_stride, float* grad_w1,
                         float* grad_b1, float* grad_gamma, float* grad_beta, float* grad_w2, float* grad_b2,
                         void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (n_points == 0) return DE_OK;
    if (!x || !grad_y || !w1_bf16 || !w1t_bf16 || !w2t_bf16 || !grad_x || !grad_w1 || !grad_w2)
        return fail(DE_EINVAL, "mlp_head_backward: NULL pointer");
    if (!de::mlp_head_supported(in_dim, hidden, out_dim))
        return fail(DE_EUNSUPPORTED, "mlp_head_backward: in_dim=%d hidden=%d out_dim=%d", in_dim, hidden, out_dim);
    if (activation < 0 || activation > 2) return fail(DE_EINVAL, "mlp_head_backward: activation=%d", activation);
    int gx_group_w = 0;
    if (gx_stride < 0) {                       // group-major grad_x: slices of W = -gx_stride floats (include/de_grid4d.h)
        gx_group_w = static_cast<int>(-gx_stride);
        if (gx_group_w < 4 || (gx_group_w & (gx_group_w - 1)) || in_dim % gx_group_w)
            return fail(DE_EINVAL, "mlp_head_backward: group width %d must be a power of two >= 4 dividing in_dim=%d", gx_group_w, in_dim);
        gx_stride = in_dim;
    }
    if (x_stride < in_dim || gx_stride < in_dim || gy_stride < out_dim) return fail(DE_EINVAL, "mlp_head_backward: row stride shorter than the row");
    const void* aligned[] = {x, grad_y, grad_x, w1_bf16, w1t_bf16, w2t_bf16};
    for (const void* q : aligned)
        if (reinterpret_cast<uintptr_t>(q) % 16) return fail(DE_EALIGN, "mlp_head_backward: pointer not 16-byte aligned");
    if (x_stride % 4 || gy_stride % 4 || gx_stride % 4) return fail(DE_EALIGN, "mlp_head_backward: row strides must be multiples of 4 floats");
    de::MlpHeadBwdParams P{};
    P.x = x; P.gy = grad_y; P.n = n_points; P.x_stride = x_stride; P.gy_stride = gy_stride;
    P.w1 = w1_bf16; P.w1t = w1t_bf16; P.w2t = w2t_bf16; P.b1 = b1; P.ln_gamma = ln_gamma; P.ln_beta = ln_beta;
    P.ln_eps = ln_eps; P.layernorm = layernorm ? 1 : 0; P.activation = activation;
    P.dx = grad_x; P.dx_stride = gx_stride; P.dx_group_w = gx_group_w; P.dw1 = grad_w1; P.db1 = grad_b1; P.dgamma = grad_gamma; P.dbeta = grad_beta;
    P.dw2 = grad_w2; P.db2 = grad_b2;
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e = de::launch_mlp_head_bwd(P, in_dim, n_sms, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "mlp_head_backward launch");
}

static int fill_field(const de_field_desc* f, const double* bbox_min, const double* bbox_span, const double* t_range,
                      de::FieldParams& Fp) {
    if (!f || !bbox_min || !bbox_span) return fail(DE_EINVAL, "field: NULL pointer");
    if (f->n_lat < 1 || f->n_lon < 1) return fail(DE_EINVAL, "field: n_lat=%d n_lon=%d must be >= 1", f->n_lat, f->n_lon);
    if (f->has_time && !t_range) return fail(DE_EINVAL, "field: time given without t_range");
    Fp = de::FieldParams{};
    Fp.lat_start = f->lat_start; Fp.lat_stop = f->lat_stop;
    Fp.lon_start = f->lon_start; Fp.lon_stop = f->lon_stop;
    Fp.lat_step = f->n_lat > 1 ? (f->lat_stop - f->lat_start) / static_cast<double>(f->n_lat - 1) : 0.0;   // np.linspace
    Fp.lon_step = f->n_lon > 1 ? (f->lon_stop - f->lon_start) / static_cast<double>(f->n_lon - 1) : 0.0;
    Fp.altitude = f->altitude; Fp.time = f->time; Fp.has_time = f->has_time ? 1 : 0;
    for (int d = 0; d < 3; ++d) { Fp.bbox_min[d] = bbox_min[d]; Fp.bbox_span[d] = bbox_span[d]; }
    Fp.t_min = t_range ? t_range[0] : 0.0;
    Fp.t_span = t_range ? t_range[1] - t_range[0] : 1.0;
    Fp.n_lat = f->n_lat; Fp.n_lon = f->n_lon;
    return DE_OK;
}

DE_API int de_field_xyzt(const de_field_desc* field_host, const double* bbox_min_host, const double* bbox_span_host,
                  const double* t_range_host, float* xyzt, void* stream) {
    de::FieldParams Fp;
    if (int rc = fill_field(field_host, bbox_min_host, bbox_span_host, t_range_host, Fp)) return rc;
    if (!xyzt) return fail(DE_EINVAL, "field_xyzt: xyzt is NULL");
    if (reinterpret_cast<uintptr_t>(xyzt) % 16) return fail(DE_EALIGN, "field_xyzt: xyzt not 16-byte aligned");
    cudaError_t e = de::launch_field_xyzt(Fp, xyzt, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "field_xyzt launch");
}

DE_API int de_grid4d_field_forward(const de_level_desc* levels_host, int n_levels, const de_field_desc* field_host,
                            const double* bbox_min_host, const double* bbox_span_host, const double* t_range_host,
                            int table_dtype, int n_features, float out_scale, float* out, int64_t out_stride,
                            float* xyzt_out, uint32_t flags, int levels_per_group, void* stream) {
    de::FieldParams Fp;
    if (int rc = fill_field(field_host, bbox_min_host, bbox_span_host, t_range_host, Fp)) return rc;
    static const float kDummy = 0.0f;
    CallArgs a{levels_host, n_levels, &kDummy, static_cast<int64_t>(Fp.n_lat) * Fp.n_lon, 4, table_dtype, n_features, flags,
               levels_per_group, true, false};
    if (int rc = validate(a)) return rc;
    if (!out) return fail(DE_EINVAL, "field_forward: out is NULL");
    if ((flags & DE_INTERP_MASK) != DE_INTERP_LINEAR) return fail(DE_EUNSUPPORTED, "field_forward: DE_INTERP_LINEAR levels only");
    if (!(n_features == 1 || n_features == 2 || n_features == 4)) return fail(DE_EUNSUPPORTED, "field_forward: n_features=%d not in {1,2,4}", n_features);
    if (n_levels > de::kMaxLaunchLevels) return fail(DE_EUNSUPPORTED, "field_forward: n_levels=%d > %d", n_levels, de::kMaxLaunchLevels);
    if (xyzt_out && reinterpret_cast<uintptr_t>(xyzt_out) % 16) return fail(DE_EALIGN, "field_forward: xyzt_out not 16-byte aligned");
    Fp.xyzt_out = xyzt_out;
    de::EncodeParams P{};
    P.out = out;
    P.n_points = a.n_points;
    P.out_stride = out_stride;
    P.coord_stride = 4;
    P.n_features = n_features;
    P.out_scale = out_scale;
    P.l2_hints = (flags & DE_FLAG_NO_L2_HINTS) ? 0 : 1;
    P.n_levels = n_levels;
    bool pow2_all = true;
    unsigned long long max_level_bytes = 1;
    for (int i = 0; i < n_levels; ++i) {
        const de_level_desc& L = levels_host[i];
        if (L.n_dims != 3) return fail(DE_EUNSUPPORTED, "field_forward: level %d has n_dims=%d (3-D grids only)", i, L.n_dims);
        if (L.mask_id) return fail(DE_EUNSUPPORTED, "field_forward: level %d carries a row mask", i);
        if ((int64_t)L.out_col + n_features > out_stride) return fail(DE_EINVAL, "level %d: out_col %u + F %d exceeds out_stride %lld", i, L.out_col, n_features, (long long)out_stride);
        fill_level(L, P.levels[i]);
        pow2_all = pow2_all && is_pow2(L.table_size);
        const unsigned long long lb = level_bytes(L, n_features, false, table_dtype);
        if (lb > max_level_bytes) max_level_bytes = lb;
    }
    int G = pick_group_size(true, n_features, levels_per_group, n_levels, max_level_bytes);
    while (G > de::field_max_group(n_features)) G >>= 1;
    P.group_size = G;
    P.n_groups = (n_levels + G - 1) / G;
    cudaError_t e = de::launch_field_forward(P, Fp, table_dtype == DE_F16, n_features, pow2_all, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "field_forward launch");
}

DE_API int de_st_embedding_forward(const float* tokens, int64_t n_tokens, int dim, const float* pos, int64_t pos_stride,
                            const float* t, const float* embed_x, const float* embed_y, int resolution,
                            const float* t_w1, const float* t_b1, const float* t_w2, const float* t_b2,
                            const float* modality, float* out, int* error_flag, void* stream) {
    if (n_tokens < 0) return fail(DE_EINVAL, "n_tokens < 0");
    if (n_tokens == 0) return DE_OK;
    if (!tokens || !out) return fail(DE_EINVAL, "st_embedding: tokens / out is NULL");
    if (dim < 4 || dim % 4) return fail(DE_EUNSUPPORTED, "st_embedding: dim=%d must be a positive multiple of 4", dim);
    if (pos && (!embed_x || !embed_y || resolution < 1)) return fail(DE_EINVAL, "st_embedding: spatial positions without tables");
    if (pos && pos_stride < 2) return fail(DE_EINVAL, "st_embedding: pos_stride=%lld < 2", (long long)pos_stride);
    if (t && (!t_w1 || !t_b1 || !t_w2 || !t_b2)) return fail(DE_EINVAL, "st_embedding: temporal positions without MLP parameters");
    de::StEmbedParams P{};
    P.tokens = tokens; P.pos = pos; P.t = t; P.ex = embed_x; P.ey = embed_y; P.w1 = t_w1; P.b1 = t_b1; P.w2 = t_w2; P.b2 = t_b2;
    P.mod = modality; P.out = out; P.err = error_flag; P.n = n_tokens; P.pos_stride = pos_stride; P.D = dim; P.R = resolution;
    cudaError_t e = de::launch_st_embed_forward(P, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "st_embedding forward launch");
}

DE_API int de_st_embedding_backward(const float* grad_out, int64_t n_tokens, int dim, const float* pos, int64_t pos_stride,
                             const float* t, int resolution, const float* t_w1, const float* t_b1, const float* t_w2,
                             float* grad_embed_x, float* grad_embed_y, float* grad_t_w1, float* grad_t_b1,
                             float* grad_t_w2, float* grad_t_b2, float* grad_modality, float* scratch, void* stream) {
    if (n_tokens < 0) return fail(DE_EINVAL, "n_tokens < 0");
    if (n_tokens == 0) return DE_OK;
    if (!grad_out) return fail(DE_EINVAL, "st_embedding_backward: grad_out is NULL");
    if (dim < 4 || dim % 4) return fail(DE_EUNSUPPORTED, "st_embedding_backward: dim=%d must be a positive multiple of 4", dim);
    if (pos && pos_stride < 2) return fail(DE_EINVAL, "st_embedding_backward: pos_stride=%lld < 2", (long long)pos_stride);
    if (pos && resolution < 1) return fail(DE_EINVAL, "st_embedding_backward: resolution=%d", resolution);
    if (pos && static_cast<size_t>(resolution) * 32 * sizeof(float) > 48 * 1024)
        return fail(DE_EUNSUPPORTED, "st_embedding_backward: resolution=%d > 384 table rows", resolution);
    if (t && (!t_w1 || !t_b1 || !t_w2 || !scratch || !grad_t_w1 || !grad_t_b1 || !grad_t_w2))
        return fail(DE_EINVAL, "st_embedding_backward: temporal branch needs w1, b1, w2, scratch and the three weight gradients");
    de::StEmbedBwdParams P{};
    P.g = grad_out; P.pos = pos; P.t = t; P.w1 = t_w1; P.b1 = t_b1; P.w2 = t_w2; P.dpre = scratch;
    P.d_ex = grad_embed_x; P.d_ey = grad_embed_y; P.d_w1 = grad_t_w1; P.d_b1 = grad_t_b1; P.d_w2 = grad_t_w2; P.d_b2 = grad_t_b2;
    P.d_mod = grad_modality; P.n = n_tokens; P.pos_stride = pos_stride; P.D = dim; P.R = resolution;
    cudaError_t e = de::launch_st_embed_backward(P, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "st_embedding backward launch");
}

static int device_sms() {
    int dev = 0, n_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    return n_sms;
}

DE_API int de_st_embedding_tc_supported(int dim) { return de::st_embed_tc_supported(dim) ? 1 : 0; }

DE_API size_t de_st_embedding_tc_workspace_bytes(int64_t n_tokens, int dim, int backward) {
    return de::st_embed_tc_workspace_bytes(n_tokens, dim, backward != 0);
}

DE_API int de_st_embedding_forward_tc(const float* tokens, int64_t n_tokens, int dim, const float* pos, int64_t pos_stride,
                               const float* t, const float* embed_x, const float* embed_y, int resolution,
                               const float* t_w1, const float* t_b1, const float* t_w2, const float* t_b2,
                               const float* modality, float* out, int* error_flag, void* workspace, size_t workspace_bytes,
                               void* stream) {
    if (n_tokens < 0) return fail(DE_EINVAL, "n_tokens < 0");
    if (n_tokens == 0) return DE_OK;
    if (!tokens || !out) return fail(DE_EINVAL, "st_embedding_tc: tokens / out is NULL");
    if (!de::st_embed_tc_supported(dim)) return fail(DE_EUNSUPPORTED, "st_embedding_tc: dim=%d must be a multiple of 128", dim);
    if (pos && (!embed_x || !embed_y || resolution < 1)) return fail(DE_EINVAL, "st_embedding_tc: spatial positions without tables");
    if (pos && pos_stride < 2) return fail(DE_EINVAL, "st_embedding_tc: pos_stride=%lld < 2", (long long)pos_stride);
    if (t && (!t_w1 || !t_b1 || !t_w2 || !t_b2)) return fail(DE_EINVAL, "st_embedding_tc: temporal positions without MLP parameters");
    if (t && (!workspace || workspace_bytes < de::st_embed_tc_workspace_bytes(n_tokens, dim, false)))
        return fail(DE_EINVAL, "st_embedding_tc: workspace of %zu bytes, %zu needed", workspace_bytes, de::st_embed_tc_workspace_bytes(n_tokens, dim, false));
    if (t && (reinterpret_cast<uintptr_t>(workspace) % 256 || reinterpret_cast<uintptr_t>(tokens) % 16 || reinterpret_cast<uintptr_t>(out) % 16 ||
              reinterpret_cast<uintptr_t>(t_w1) % 16 || reinterpret_cast<uintptr_t>(t_b1) % 16))
        return fail(DE_EALIGN, "st_embedding_tc: workspace must be 256-byte, tokens / out / t_w1 / t_b1 16-byte aligned");
    if (n_tokens > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "st_embedding_tc: n_tokens too large");
    de::StEmbedParams P{};
    P.tokens = tokens; P.pos = pos; P.t = t; P.ex = embed_x; P.ey = embed_y; P.w1 = t_w1; P.b1 = t_b1; P.w2 = t_w2; P.b2 = t_b2;
    P.mod = modality; P.out = out; P.err = error_flag; P.n = n_tokens; P.pos_stride = pos_stride; P.D = dim; P.R = resolution;
    cudaError_t e = de::launch_st_embed_forward_tc(P, workspace, device_sms(), static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "st_embedding_tc forward launch");
}

DE_API int de_st_embedding_backward_tc(const float* grad_out, int64_t n_tokens, int dim, const float* pos, int64_t pos_stride,
                                const float* t, int resolution, const float* t_w1, const float* t_b1, const float* t_w2,
                                float* grad_embed_x, float* grad_embed_y, float* grad_t_w1, float* grad_t_b1,
                                float* grad_t_w2, float* grad_t_b2, float* grad_modality, void* workspace, size_t workspace_bytes,
                                void* stream) {
    if (n_tokens < 0) return fail(DE_EINVAL, "n_tokens < 0");
    if (n_tokens == 0) return DE_OK;
    if (!grad_out) return fail(DE_EINVAL, "st_embedding_backward_tc: grad_out is NULL");
    if (!de::st_embed_tc_supported(dim)) return fail(DE_EUNSUPPORTED, "st_embedding_backward_tc: dim=%d must be a multiple of 128", dim);
    if (pos && pos_stride < 2) return fail(DE_EINVAL, "st_embedding_backward_tc: pos_stride=%lld < 2", (long long)pos_stride);
    if (pos && resolution < 1) return fail(DE_EINVAL, "st_embedding_backward_tc: resolution=%d", resolution);
    if (pos && static_cast<size_t>(resolution) * 32 * sizeof(float) > 48 * 1024)
        return fail(DE_EUNSUPPORTED, "st_embedding_backward_tc: resolution=%d > 384 table rows", resolution);
    if (t && (!t_w1 || !t_b1 || !t_w2 || !grad_t_w1 || !grad_t_b1 || !grad_t_w2))
        return fail(DE_EINVAL, "st_embedding_backward_tc: temporal branch needs w1, b1, w2 and the three weight gradients");
    if (t && (!workspace || workspace_bytes < de::st_embed_tc_workspace_bytes(n_tokens, dim, true)))
        return fail(DE_EINVAL, "st_embedding_backward_tc: workspace of %zu bytes, %zu needed", workspace_bytes, de::st_embed_tc_workspace_bytes(n_tokens, dim, true));
    if (t && (reinterpret_cast<uintptr_t>(workspace) % 256 || reinterpret_cast<uintptr_t>(grad_t_w2) % 16 ||
              reinterpret_cast<uintptr_t>(t_w1) % 16 || reinterpret_cast<uintptr_t>(t_b1) % 16))
        return fail(DE_EALIGN, "st_embedding_backward_tc: workspace must be 256-byte, grad_t_w2 / t_w1 / t_b1 16-byte aligned");
    if (n_tokens > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "st_embedding_backward_tc: n_tokens too large");
    de::StEmbedBwdParams P{};
    P.g = grad_out; P.pos = pos; P.t = t; P.w1 = t_w1; P.b1 = t_b1; P.w2 = t_w2; P.dpre = nullptr;
    P.d_ex = grad_embed_x; P.d_ey = grad_embed_y; P.d_w1 = grad_t_w1; P.d_b1 = grad_t_b1; P.d_w2 = grad_t_w2; P.d_b2 = grad_t_b2;
    P.d_mod = grad_modality; P.n = n_tokens; P.pos_stride = pos_stride; P.D = dim; P.R = resolution;
    cudaError_t e = de::launch_st_embed_backward_tc(P, workspace, device_sms(), static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "st_embedding_tc backward launch");
}

DE_API int de_gemm_bf16(const void* a_bf16, int64_t lda, const void* b_bf16, int64_t ldb, int64_t m, int n, int64_t k,
                 const float* bias, void* c, int64_t ldc, int out_dtype, int transposed, int accumulate, void* stream) {
    if (m < 0 || n < 1 || k < 1) return fail(DE_EINVAL, "gemm_bf16: m=%lld n=%d k=%lld", (long long)m, n, (long long)k);
    if (m == 0) return DE_OK;
    if (!a_bf16 || !b_bf16 || !c) return fail(DE_EINVAL, "gemm_bf16: NULL pointer");
    if (out_dtype != DE_F32 && out_dtype != DE_BF16) return fail(DE_EUNSUPPORTED, "gemm_bf16: out_dtype=%d", out_dtype);
    if (accumulate && (out_dtype != DE_F32 || bias)) return fail(DE_EUNSUPPORTED, "gemm_bf16: accumulation is fp32 and takes no bias");
    if (reinterpret_cast<uintptr_t>(a_bf16) % 16 || reinterpret_cast<uintptr_t>(b_bf16) % 16 || lda % 8 || ldb % 8)
        return fail(DE_EALIGN, "gemm_bf16: operand rows must be 16-byte aligned (TMA)");
    if (transposed ? (lda < m || ldb < n) : (lda < k || ldb < k)) return fail(DE_EINVAL, "gemm_bf16: leading dimension shorter than the row");
    if (ldc < n) return fail(DE_EINVAL, "gemm_bf16: ldc shorter than the row");
    if (m > 0x7fffffffll || k > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "gemm_bf16: m=%lld k=%lld too large", (long long)m, (long long)k);
    de::GemmParams P{};
    P.M = m; P.N = n; P.K = k; P.c = c; P.ldc = ldc; P.bias = bias;
    cudaError_t e = de::launch_gemm_bf16(a_bf16, lda, b_bf16, ldb, transposed ? 1 : 0, P, accumulate ? 2 : (out_dtype == DE_BF16 ? 1 : 0),
                                         device_sms(), static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "gemm_bf16 launch (TMA descriptor or kernel)");
}

DE_API int de_projection_supported(int in_dim, int hidden, int out_dim) {
    return (de::proj_hidden_supported(in_dim, hidden) && out_dim >= 8 && out_dim % 8 == 0) ? 1 : 0;
}

DE_API int de_projection_forward(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                          const void* w1_bf16, const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps,
                          const void* w2_bf16, const float* b2, int hidden, int out_dim,
                          void* hidden_bf16, float* row_stats, float* y, int64_t y_stride, void* stream) {
    return de_projection_forward_ex(x, n_points, in_dim, x_stride, w1_bf16, b1, ln_gamma, ln_beta, ln_eps, w2_bf16, b2, hidden, out_dim,
                                    hidden_bf16, row_stats, y, DE_F32, y_stride, stream);
}

DE_API int de_projection_forward_ex(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                             const void* w1_bf16, const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps,
                             const void* w2_bf16, const float* b2, int hidden, int out_dim,
                             void* hidden_bf16, float* row_stats, void* y, int y_dtype, int64_t y_stride, void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (y_dtype != DE_F32 && y_dtype != DE_BF16) return fail(DE_EUNSUPPORTED, "projection: y_dtype %d (DE_F32 or DE_BF16)", y_dtype);
    if (n_points == 0) return DE_OK;
    if (!x || !w1_bf16 || !b1 || !ln_gamma || !ln_beta || !w2_bf16 || !hidden_bf16 || !y) return fail(DE_EINVAL, "projection: NULL pointer");
    if (!de_projection_supported(in_dim, hidden, out_dim))
        return fail(DE_EUNSUPPORTED, "projection: in_dim=%d hidden=%d out_dim=%d (in_dim in {32,48,64,96,128}, hidden a multiple of 128, out_dim of 8)", in_dim, hidden, out_dim);
    if (x_stride < in_dim || y_stride < out_dim) return fail(DE_EINVAL, "projection: row stride shorter than the row");
    if (reinterpret_cast<uintptr_t>(x) % 16 || x_stride % 4) return fail(DE_EALIGN, "projection: x rows not 16-byte aligned");
    const void* aligned[] = {w1_bf16, w2_bf16, hidden_bf16, y, b1, ln_gamma, ln_beta};
    for (const void* q : aligned)
        if (reinterpret_cast<uintptr_t>(q) % 16) return fail(DE_EALIGN, "projection: pointer not 16-byte aligned");
    if (n_points > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "projection: n_points too large");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int n_sms = device_sms();
    de::ProjHiddenParams H{};
    H.x = x; H.n = n_points; H.x_stride = x_stride; H.w1 = w1_bf16; H.b1 = b1; H.ln_gamma = ln_gamma; H.ln_beta = ln_beta;
    H.ln_eps = ln_eps; H.hidden = hidden; H.hidden_out = hidden_bf16; H.hidden_stride = hidden; H.stats_out = row_stats;
    cudaError_t e = de::launch_proj_hidden(H, in_dim, n_sms, s);
    if (e != cudaSuccess) return cuda_fail(e, "projection hidden launch");
    de::GemmParams G{};
    G.M = n_points; G.N = out_dim; G.K = hidden; G.c = y; G.ldc = y_stride; G.bias = b2;
    e = de::launch_gemm_bf16(hidden_bf16, hidden, w2_bf16, hidden, 0, G, y_dtype == DE_BF16 ? 1 : 0, n_sms, s);
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "projection GEMM launch");
}

static size_t align256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

DE_API size_t de_projection_backward_workspace_bytes(int64_t n_points, int in_dim, int hidden, int out_dim) {
    if (n_points <= 0 || in_dim <= 0 || hidden <= 0 || out_dim <= 0) return 0;
    const size_t n = static_cast<size_t>(n_points);
    const size_t nt = (n + 7) & ~static_cast<size_t>(7);      // leading dimension of the transposed copies (fallback path)
    return align256(n * out_dim * 2) + align256(n * hidden * 2) + align256(n * in_dim * 2) +
           align256(nt * out_dim * 2) + align256(nt * hidden * 2);
}

DE_API int de_projection_backward(const float* x, const float* grad_y, int64_t n_points, int in_dim, int64_t x_stride,
                           int64_t gy_stride, const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                           const float* b1, const float* ln_gamma, const float* ln_beta, int hidden, int out_dim,
                           const void* hidden_bf16, const float* row_stats,
                           float* grad_x, int64_t gx_stride, float* grad_w1, float* grad_b1, float* grad_gamma,
                           float* grad_beta, float* grad_w2, float* grad_b2,
                           void* workspace, size_t workspace_bytes, void* stream) {
    return de_projection_backward_ex(x, grad_y, DE_F32, n_points, in_dim, x_stride, gy_stride, w1_bf16, w1t_bf16, w2t_bf16, b1, ln_gamma,
                                     ln_beta, hidden, out_dim, hidden_bf16, row_stats, grad_x, gx_stride, grad_w1, grad_b1, grad_gamma,
                                     grad_beta, grad_w2, grad_b2, workspace, workspace_bytes, stream);
}

DE_API int de_projection_backward_ex(const float* x, const void* grad_y, int gy_dtype, int64_t n_points, int in_dim, int64_t x_stride,
                              int64_t gy_stride, const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                              const float* b1, const float* ln_gamma, const float* ln_beta, int hidden, int out_dim,
                              const void* hidden_bf16, const float* row_stats,
                              float* grad_x, int64_t gx_stride, float* grad_w1, float* grad_b1, float* grad_gamma,
                              float* grad_beta, float* grad_w2, float* grad_b2,
                              void* workspace, size_t workspace_bytes, void* stream) {
    if (n_points < 0) return fail(DE_EINVAL, "n_points < 0");
    if (gy_dtype != DE_F32 && gy_dtype != DE_BF16) return fail(DE_EUNSUPPORTED, "projection_backward: gy_dtype %d (DE_F32 or DE_BF16)", gy_dtype);
    if (n_points == 0) return DE_OK;
    if (!x || !grad_y || !w1_bf16 || !w1t_bf16 || !w2t_bf16 || !b1 || !ln_gamma || !ln_beta || !hidden_bf16 || !row_stats ||
        !grad_w1 || !grad_w2 || !workspace)
        return fail(DE_EINVAL, "projection_backward: NULL pointer");
    if (!de_projection_supported(in_dim, hidden, out_dim)) return fail(DE_EUNSUPPORTED, "projection_backward: in_dim=%d hidden=%d out_dim=%d", in_dim, hidden, out_dim);
    if (workspace_bytes < de_projection_backward_workspace_bytes(n_points, in_dim, hidden, out_dim))
        return fail(DE_EINVAL, "projection_backward: workspace of %zu bytes is too small", workspace_bytes);
    if (x_stride < in_dim || gy_stride < out_dim || (grad_x && gx_stride < in_dim)) return fail(DE_EINVAL, "projection_backward: row stride shorter than the row");
    const void* aligned[] = {x, grad_y, w1_bf16, w1t_bf16, w2t_bf16, hidden_bf16, workspace, grad_x, b1, ln_gamma, ln_beta};
    for (const void* q : aligned)
        if (reinterpret_cast<uintptr_t>(q) % 16) return fail(DE_EALIGN, "projection_backward: pointer not 16-byte aligned");
    if (x_stride % 4 || gy_stride % (gy_dtype == DE_BF16 ? 8 : 4))
        return fail(DE_EALIGN, "projection_backward: row strides must be multiples of 16 bytes");
    if (n_points > 0x7fffffffll) return fail(DE_EUNSUPPORTED, "projection_backward: n_points too large");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int n_sms = device_sms();
    const size_t n = static_cast<size_t>(n_points);
    char* ws = static_cast<char*>(workspace);
    const void* gy16 = gy_dtype == DE_BF16 ? grad_y : ws;      // bf16 gradients are consumed where they lie (TMA: any 16-byte row pitch)
    const int64_t gy16_ld = gy_dtype == DE_BF16 ? gy_stride : out_dim;
    void* da = ws + align256(n * out_dim * 2);
    void* x16 = ws + align256(n * out_dim * 2) + align256(n * hidden * 2);
    // 1. bf16 copy of grad_y (fp32 input only), db2 = column sums
    cudaError_t e = cudaSuccess;
    if (gy_dtype == DE_BF16) {
        if (grad_b2) e = de::launch_colsum_bf16(grad_y, gy_stride, n_points, out_dim, grad_b2, n_sms, s);
    } else {
        e = de::launch_cast_colsum(static_cast<const float*>(grad_y), gy_stride, ws, out_dim, n_points, out_dim, grad_b2, n_sms, s);
    }
    if (e != cudaSuccess) return cuda_fail(e, "projection_backward: cast / column-sum launch");
    // 2. da (n, hidden) = gy (n, out) . W2 (out, hidden)
    {
        de::GemmParams G{};
        G.M = n_points; G.N = hidden; G.K = out_dim; G.c = da; G.ldc = hidden;
        e = de::launch_gemm_bf16(gy16, gy16_ld, w2t_bf16, out_dim, 0, G, 1, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: da GEMM launch");
    }
    // 3. dpre in place of da; dgamma, dbeta, db1; bf16 copy of x
    {
        de::ProjDpreParams D{};
        D.x = x; D.n = n_points; D.x_stride = x_stride; D.w1 = w1_bf16; D.b1 = b1; D.ln_gamma = ln_gamma; D.ln_beta = ln_beta;
        D.stats = row_stats; D.dact = da; D.dact_stride = hidden; D.dpre_out = da; D.x_bf16_out = x16;
        D.dgamma = grad_gamma; D.dbeta = grad_beta; D.db1 = grad_b1; D.hidden = hidden;
        e = de::launch_proj_dpre(D, in_dim, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dpre launch");
    }
    // 4. dW2 (out, hidden) += gy^T . a ;  5. dW1 (hidden, in) += dpre^T . x   (contraction over the points)
    if (DE_KNOB("DE_PROJ_DW_TRANSPOSE", 0, 0, 1)) {
        // fallback operand path: explicit transposes, then the K-major GEMM (K = points) -- same results, ~2x the bytes
        const size_t nt = (n + 7) & ~static_cast<size_t>(7);
        char* t0 = ws + align256(n * out_dim * 2) + align256(n * hidden * 2) + align256(n * in_dim * 2);
        char* t1 = t0 + align256(nt * out_dim * 2);
        e = de::launch_transpose_bf16(gy16, gy16_ld, t0, nt, n_points, out_dim, s);
        if (e == cudaSuccess) e = de::launch_transpose_bf16(hidden_bf16, hidden, t1, nt, n_points, hidden, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: transpose launch");
        de::GemmParams G{};
        G.M = out_dim; G.N = hidden; G.K = n_points; G.c = grad_w2; G.ldc = hidden;
        e = de::launch_gemm_bf16(t0, nt, t1, nt, 0, G, 2, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dW2 GEMM launch");
        e = de::launch_transpose_bf16(da, hidden, t1, nt, n_points, hidden, s);
        if (e == cudaSuccess) e = de::launch_transpose_bf16(x16, in_dim, t0, nt, n_points, in_dim, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: transpose launch");
        G = de::GemmParams{};
        G.M = hidden; G.N = in_dim; G.K = n_points; G.c = grad_w1; G.ldc = in_dim;
        e = de::launch_gemm_bf16(t1, nt, t0, nt, 0, G, 2, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dW1 GEMM launch");
    } else {
        de::GemmParams G{};
        G.M = out_dim; G.N = hidden; G.K = n_points; G.c = grad_w2; G.ldc = hidden;
        e = de::launch_gemm_bf16(gy16, gy16_ld, hidden_bf16, hidden, 1, G, 2, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dW2 GEMM launch");
        G = de::GemmParams{};
        G.M = hidden; G.N = in_dim; G.K = n_points; G.c = grad_w1; G.ldc = in_dim;
        e = de::launch_gemm_bf16(da, hidden, x16, in_dim, 1, G, 2, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dW1 GEMM launch");
    }
    // 6. dx (n, in) = dpre (n, hidden) . W1 (hidden, in)
    if (grad_x) {
        de::GemmParams G{};
        G.M = n_points; G.N = in_dim; G.K = hidden; G.c = grad_x; G.ldc = gx_stride;
        e = de::launch_gemm_bf16(da, hidden, w1t_bf16, hidden, 0, G, 0, n_sms, s);
        if (e != cudaSuccess) return cuda_fail(e, "projection_backward: dx GEMM launch");
    }
    return DE_OK;
}

DE_API int de_set_l2_fetch_granularity(int bytes) {
    if (bytes != 32 && bytes != 64 && bytes != 128) return fail(DE_EINVAL, "l2 fetch granularity must be 32, 64 or 128, got %d", bytes);
    cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, static_cast<size_t>(bytes));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "cudaDeviceSetLimit(MaxL2FetchGranularity)");
}

DE_API int de_get_l2_fetch_granularity(int* bytes) {
    if (!bytes) return fail(DE_EINVAL, "bytes is NULL");
    size_t v = 0;
    cudaError_t e = cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetLimit(MaxL2FetchGranularity)");
    *bytes = static_cast<int>(v);
    return DE_OK;
}

DE_API int de_set_persisting_l2(long long bytes) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    int max_bytes = 0;
    e = cudaDeviceGetAttribute(&max_bytes, cudaDevAttrMaxPersistingL2CacheSize, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute(MaxPersistingL2CacheSize)");
    if (bytes < 0 || bytes > max_bytes) bytes = max_bytes;
    e = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, static_cast<size_t>(bytes));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "cudaDeviceSetLimit(PersistingL2CacheSize)");
}

DE_API int de_get_persisting_l2(long long* bytes, long long* max_bytes) {
    if (!bytes || !max_bytes) return fail(DE_EINVAL, "NULL pointer");
    int dev = 0, mx = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    e = cudaDeviceGetAttribute(&mx, cudaDevAttrMaxPersistingL2CacheSize, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute(MaxPersistingL2CacheSize)");
    size_t v = 0;
    e = cudaDeviceGetLimit(&v, cudaLimitPersistingL2CacheSize);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetLimit(PersistingL2CacheSize)");
    *bytes = static_cast<long long>(v);
    *max_bytes = mx;
    return DE_OK;
}

DE_API int de_sum_f32(const float* x, int64_t n, double* result, void* stream) {
    if (n < 0 || (n > 0 && (!x || !result))) return fail(DE_EINVAL, "sum_f32: bad arguments");
    cudaError_t e = de::launch_sum_f32(x, n, result, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "sum_f32 launch");
}

DE_API int de_probe_gather(void* table, int64_t n_entries, int entry_bytes, int mode, int64_t n_threads,
                    int per_thread, int pair_pattern, void* sink, void* stream) {
    if (!table || !sink) return fail(DE_EINVAL, "probe: NULL pointer");
    if (n_entries <= 0 || (n_entries & (n_entries - 1)) || n_entries > (1ll << 32)) return fail(DE_EINVAL, "probe: n_entries must be a power of two <= 2^32");
    if (entry_bytes != 4 && entry_bytes != 8 && entry_bytes != 16) return fail(DE_EUNSUPPORTED, "probe: entry_bytes=%d", entry_bytes);
    if (mode != 0 && mode != 1) return fail(DE_EINVAL, "probe: mode=%d", mode);
    if (pair_pattern != 0 && pair_pattern != 1 && pair_pattern != 2 && pair_pattern != 4 && pair_pattern != 8) return fail(DE_EINVAL, "probe: pair_pattern=%d", pair_pattern);
    if (n_threads <= 0 || per_thread <= 0 || per_thread % 8) return fail(DE_EINVAL, "probe: per_thread must be a positive multiple of 8");
    cudaError_t e = de::launch_probe(table, n_entries, entry_bytes, mode, n_threads, per_thread, pair_pattern, sink, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? DE_OK : cuda_fail(e, "probe launch");
}

}  // extern "C"
