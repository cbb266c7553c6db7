// One host process driving several GPUs through the C-ABI alone ("hyb_ctx_create(ndev, dev[])",
// "one stream per device ... created in ctx_create"): what a Julia host needs to use a whole node with
// nothing but ccalls.
//
// Shape (north_star): the cones are partitioned by block across the GPUs; every GPU builds the
// partial Schur matrix of its cones, sum_k GQ2_k' H~_k GQ2_k (qrchol.jl:226-254), and the FP64 DMMA syrk's
// epilogue stores every finished tile straight into that GPU's slot in the FACTORIZING GPU's memory over
// NVLink (peer access inside one process -- no IPC handles, no staging copy, no collective library).  The
// root adds the slots in device order (fixed order => bitwise reproducible) and factorizes.  Everything
// else -- vector oracles, line search, Cholesky, direction solves -- is the root's single-GPU context,
// reachable through hyb_multi_root(), so all other entry points apply unchanged.
//
// Host threading: the per-device work of one update is issued from one short-lived host thread per
// device (the existing entry points synchronise at return, so sequential calls from one thread would
// serialise the GPUs).
#include <chrono>
#include <thread>

#include "ctx.cuh"

struct hyb_multi {
    int ndev = 0;
    std::vector<int> devices;
    hyb_ctx* root = nullptr;
    std::vector<hyb_ctx*> shard;         // shard[d] lives on devices[d]; shard[0] shares the root's G*Q2 rows
    std::vector<int64_t> row_lo, row_hi; // rows of s / z owned by shard d
    std::vector<double> t_asm;
    // reduce-scatter of the partial Schur matrices: device o owns the tile columns j with j % ndev == o and holds
    // ndev compact slots (one per source shard) of slot_elems doubles for them
    std::vector<double*> own_slots;
    int64_t slot_elems = 0;
    bool loaded = false;
};

namespace {

// Owner `o` adds the ndev partial copies of its tile columns in source order (deterministic) and stores the sums
// into the factorizing GPU's Schur matrix S (peer memory unless o is that GPU).  Compact column block jb of the
// slots is global column block jb * nown + o; upper triangle only.
__global__ void sum_owned_kernel(const double* __restrict__ slots, int nsrc, int64_t slot_elems, int64_t N, int64_t ld, int o,
                                 int nown, double* __restrict__ S) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t cc = blockIdx.y;  // compact column
    const int64_t jb = cc >> 7, gcol = ((jb * nown + o) << 7) + (cc & 127);
    if (gcol >= N || r > gcol) return;
    const int64_t off = r + cc * ld;
    double acc = 0.0;
    for (int k = 0; k < nsrc; ++k) acc += slots[k * slot_elems + off];
    S[r + gcol * ld] = acc;
}

// Contiguous partition of the cones into ndev row ranges of about q / ndev rows each.  Cones are atoms,
// except Nonnegative cones: the orthant is a product of 1-dimensional cones, so it is cut exactly at the
// target (on an even row, which keeps the shard's operands TMA-aligned) and LPs shard too.
struct ShardPlan {
    std::vector<std::vector<hyb_cone_desc>> descs;  // per shard, offsets rebased to the shard
    std::vector<int64_t> lo, hi;
};

ShardPlan plan_shards(int64_t q, int64_t ncones, const hyb_cone_desc* descs, int ndev) {
    ShardPlan p;
    p.descs.resize(ndev);
    p.lo.assign(ndev, 0);
    p.hi.assign(ndev, 0);
    int d = 0;
    int64_t start = 0;  // first row of shard d
    auto target_end = [&](int dd) { return (dd == ndev - 1) ? q : ((q * (dd + 1)) / ndev) & ~(int64_t)1; };
    auto close = [&](int64_t row) {
        p.lo[d] = start;
        p.hi[d] = row;
        start = row;
        ++d;
    };
    for (int64_t k = 0; k < ncones; ++k) {
        hyb_cone_desc c = descs[k];
        int64_t off = c.offset, left = c.dim;
        while (left > 0) {
            const int64_t tgt = target_end(d);
            if (c.type == HYB_CONE_NONNEGATIVE && d < ndev - 1 && off < tgt && off + left > tgt) {
                hyb_cone_desc part = c;  // head of the orthant stays in shard d
                part.offset = off - start;
                part.dim = tgt - off;
                p.descs[d].push_back(part);
                left -= part.dim;
                off += part.dim;
                close(tgt);
                continue;
            }
            hyb_cone_desc part = c;
            part.offset = off - start;
            part.dim = left;
            p.descs[d].push_back(part);
            off += left;
            left = 0;
            // an atom that reaches past the target ends the shard (the last shard takes the rest)
            while (d < ndev - 1 && off >= target_end(d)) close(off);
        }
    }
    while (d < ndev) close(q);
    return p;
}

}  // namespace

extern "C" {

void hyb_multi_destroy(hyb_multi* m) {
    if (!m) return;
    for (size_t o = 0; o < m->own_slots.size(); ++o)
        if (m->own_slots[o]) {
            cudaSetDevice(m->devices[o]);
            cudaFree(m->own_slots[o]);
        }
    for (hyb_ctx* s : m->shard) hyb_ctx_destroy(s);
    hyb_ctx_destroy(m->root);
    delete m;
}

hyb_multi* hyb_multi_create(int ndev, const int* devices) {
    if (ndev < 1 || !devices) {
        hyb_set_error("hyb_multi_create: need at least one device");
        return nullptr;
    }
    hyb_multi* m = new hyb_multi();
    m->ndev = ndev;
    m->devices.assign(devices, devices + ndev);
    m->root = hyb_ctx_create(devices[0]);
    if (!m->root) {
        delete m;
        return nullptr;
    }
    for (int d = 0; d < ndev; ++d) {
        hyb_ctx* s = hyb_ctx_create(devices[d]);
        if (!s) {
            hyb_multi_destroy(m);
            return nullptr;
        }
        s->want_assemble_only = true;
        m->shard.push_back(s);
        // kernels on device d store tiles into every other device's memory (reduce-scatter of the partial Schur
        // matrices) and the owners store their sums into the first device's
        for (int e2 = 0; e2 < ndev; ++e2) {
            if (devices[e2] == devices[d]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[d], devices[e2]);
            if (!can) {
                hyb_set_error("hyb_multi_create: device %d cannot access the memory of device %d (no P2P / NVLink path)", devices[d],
                              devices[e2]);
                hyb_multi_destroy(m);
                return nullptr;
            }
            cudaSetDevice(devices[d]);
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[e2], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                hyb_set_error("hyb_multi_create: cudaDeviceEnablePeerAccess(%d -> %d): %s", devices[d], devices[e2],
                              cudaGetErrorString(e));
                hyb_multi_destroy(m);
                return nullptr;
            }
            cudaGetLastError();
        }
    }
    m->row_lo.assign(ndev, 0);
    m->row_hi.assign(ndev, 0);
    m->t_asm.assign(ndev, 0.0);
    return m;
}

hyb_ctx* hyb_multi_root(hyb_multi* m) { return m ? m->root : nullptr; }

// The partition hyb_multi_model_load would use, without touching a GPU: row ranges per device and the number of
// (possibly split) cones each device gets.  Pure host code -- lets a caller size its buffers, and lets the CPU
// test-suite check the partition rules.
int hyb_multi_plan(int64_t q, int64_t ncones, const hyb_cone_desc* descs, int ndev, int64_t* row_lo, int64_t* row_hi,
                   int64_t* cones_per_dev) {
    if (ndev < 1 || (ncones > 0 && !descs)) {
        hyb_set_error("hyb_multi_plan: bad arguments");
        return -3;
    }
    const ShardPlan plan = plan_shards(q, ncones, descs, ndev);
    for (int d = 0; d < ndev; ++d) {
        if (row_lo) row_lo[d] = plan.lo[d];
        if (row_hi) row_hi[d] = plan.hi[d];
        if (cones_per_dev) cones_per_dev[d] = (int64_t)plan.descs[d].size();
    }
    return 0;
}

int hyb_multi_shard_rows(hyb_multi* m, int64_t* row_lo, int64_t* row_hi) {
    if (!m || !m->loaded) return -1;
    for (int d = 0; d < m->ndev; ++d) {
        if (row_lo) row_lo[d] = m->row_lo[d];
        if (row_hi) row_hi[d] = m->row_hi[d];
    }
    return 0;
}

int hyb_multi_model_load(hyb_multi* m, int64_t q, int64_t N, const double* GQ2, int64_t ncones, const hyb_cone_desc* descs) {
    if (!m) return -1;
    m->loaded = false;
    HYB_CHECK(hyb_model_load(m->root, q, N, GQ2, ncones, descs));
    if (N == 0) {  // cones-only model (initial-point oracles, Solvers.jl:749-767): nothing to shard
        m->loaded = true;
        for (int d = 0; d < m->ndev; ++d) m->row_lo[d] = m->row_hi[d] = 0;
        return 0;
    }
    // owner slots: ceil(T / ndev) column blocks of 128 columns x ldS rows per (owner, source)
    const int64_t T = (N + 127) / 128, ncb = (T + m->ndev - 1) / m->ndev;
    m->slot_elems = ncb * 128 * m->root->ldS;
    for (size_t o = 0; o < m->own_slots.size(); ++o)
        if (m->own_slots[o]) {
            cudaSetDevice(m->devices[o]);
            cudaFree(m->own_slots[o]);
        }
    m->own_slots.assign(m->ndev, nullptr);
    for (int o = 0; o < m->ndev; ++o) {
        HYB_CUDA(cudaSetDevice(m->devices[o]));
        HYB_CUDA(cudaMalloc(&m->own_slots[o], sizeof(double) * m->slot_elems * m->ndev));
        HYB_CUDA(cudaMemset(m->own_slots[o], 0, sizeof(double) * m->slot_elems * m->ndev));
    }
    const ShardPlan plan = plan_shards(q, ncones, descs, m->ndev);
    for (int d = 0; d < m->ndev; ++d) {
        hyb_ctx* s = m->shard[d];
        const int64_t lo = plan.lo[d], rows = plan.hi[d] - plan.lo[d];
        m->row_lo[d] = lo;
        m->row_hi[d] = plan.hi[d];
        const std::vector<hyb_cone_desc>& ds = plan.descs[d];
        if (m->devices[d] == m->devices[0])
            HYB_CHECK(hyb_model_load_impl(s, rows, N, nullptr, q, (int64_t)ds.size(), ds.data(), m->root, lo));
        else
            HYB_CHECK(hyb_model_load_impl(s, rows, N, GQ2 + lo, q, (int64_t)ds.size(), ds.data(), nullptr, 0));
        // where this shard's tile columns go: its slot in every owner's memory (same process: plain peer pointers)
        s->own_bases_host.resize(m->ndev);
        for (int o = 0; o < m->ndev; ++o) s->own_bases_host[o] = m->own_slots[o] + (int64_t)d * m->slot_elems;
        s->own_slot_elems = m->slot_elems;
        HYB_CUDA(cudaSetDevice(s->device));
        cudaFree(s->own_bases_dev);
        HYB_CUDA(cudaMalloc(&s->own_bases_dev, sizeof(double*) * m->ndev));
        HYB_CUDA(cudaMemcpy(s->own_bases_dev, s->own_bases_host.data(), sizeof(double*) * m->ndev, cudaMemcpyHostToDevice));
        s->peer_slot = nullptr;
        s->peer_world = m->ndev;
        s->peer_rank = d;
    }
    m->loaded = true;
    return 0;
}

int hyb_multi_schur_update(hyb_multi* m, double* t_assemble_ms, double* t_reduce_ms, double* t_fact_ms, int* fact_kind) {
    if (!m || !m->loaded) return -1;
    hyb_ctx* r = m->root;
    if (t_assemble_ms) *t_assemble_ms = 0;
    if (t_reduce_ms) *t_reduce_ms = 0;
    if (r->N == 0) return hyb_schur_factor(r, t_fact_ms, fact_kind);
    if (!r->point_loaded) {
        hyb_set_error("hyb_multi_schur_update: no point loaded in the root context");
        return -3;
    }
    HYB_CUDA(cudaSetDevice(r->device));
    // HYB_MULTI_TRACE=1: host wall-clock stamps of this call's phases on stderr (a measurement aid)
    static const bool trace = getenv("HYB_MULTI_TRACE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    HYB_CUDA(cudaStreamSynchronize(r->st));
    const double t_sync = since();
    std::vector<double> t_copy(m->ndev, 0.0), t_upd(m->ndev, 0.0), t_done(m->ndev, 0.0);
    std::vector<int> rc(m->ndev, 0);
    std::vector<std::string> err(m->ndev);
    const bool root_lazy = r->hess_fact_lazy;
    r->hess_fact_lazy = false;
    auto work = [&](int d) {
        hyb_ctx* s = m->shard[d];
        const int64_t lo = m->row_lo[d], rows = m->row_hi[d] - lo;
        m->t_asm[d] = 0.0;
        auto fail = [&](int code) {
            rc[d] = code;
            err[d] = hyb_get_error();
        };
        if (cudaSetDevice(s->device) != cudaSuccess) {
            hyb_set_error("cudaSetDevice(%d) failed", s->device);
            return fail(-1);
        }
        if (rows > 0) {
            // this shard's slice of the accepted point (load_point / load_dual_point, Cones.jl:164-170)
            cudaError_t e = cudaMemcpyPeerAsync(s->in_primal, s->device, r->in_primal + lo, r->device, sizeof(double) * rows, s->st);
            if (e == cudaSuccess)
                e = cudaMemcpyPeerAsync(s->in_dual, s->device, r->in_dual + lo, r->device, sizeof(double) * rows, s->st);
            if (e != cudaSuccess) {
                hyb_set_error("hyb_multi_schur_update: peer copy to device %d: %s", s->device, cudaGetErrorString(e));
                return fail(-1);
            }
        }
        s->loaded_scal = r->loaded_scal;
        s->point_loaded = true;
        s->updated_irtmu = NAN;
        s->hess_fact_lazy = root_lazy;  // the shards take the branch the full model would (Cones.jl:193-199)
        t_copy[d] = since();
        int c = hyb_ensure_updated(s, s->loaded_scal);
        t_upd[d] = since();
        if (c == 0) c = hyb_schur_assemble_into(s, nullptr, &m->t_asm[d]);  // tiles go straight to their owners
        t_done[d] = since();
        if (c != 0) fail(c);
    };
    std::vector<std::thread> th;
    for (int d = 1; d < m->ndev; ++d) th.emplace_back(work, d);
    work(0);
    for (auto& t : th) t.join();
    const double t_join = since();
    for (int d = 0; d < m->ndev; ++d)
        if (rc[d] != 0) {
            hyb_set_error("device %d: %s", m->devices[d], err[d].c_str());
            return rc[d];
        }
    double ta = 0.0;
    for (int d = 0; d < m->ndev; ++d) ta = ta > m->t_asm[d] ? ta : m->t_asm[d];
    if (t_assemble_ms) *t_assemble_ms = ta;
    // every owner sums its tile columns and stores them into the root's Schur matrix
    {
        const double t_before = since();
        const int64_t N = r->N, ncols = (m->slot_elems / r->ldS);
        for (int o = 0; o < m->ndev; ++o) {
            hyb_ctx* s = m->shard[o];
            HYB_CUDA(cudaSetDevice(s->device));
            dim3 grid((unsigned)((N + 255) / 256), (unsigned)ncols);
            sum_owned_kernel<<<grid, 256, 0, s->st>>>(m->own_slots[o], m->ndev, m->slot_elems, N, r->ldS, o, m->ndev, r->S);
            HYB_CUDA(cudaGetLastError());
            ++g_misc_launches;
        }
        for (int o = 0; o < m->ndev; ++o) {
            HYB_CUDA(cudaSetDevice(m->shard[o]->device));
            HYB_CUDA(cudaStreamSynchronize(m->shard[o]->st));
        }
        HYB_CUDA(cudaSetDevice(r->device));
        r->fact_ok = false;
        if (t_reduce_ms) *t_reduce_ms = since() - t_before;
        r->dev_ms += since() - t_before;
    }
    const double t_sum = since();
    const int rcf = hyb_schur_factor(r, t_fact_ms, fact_kind);
    if (trace) {
        fprintf(stderr, "multi trace (ms): sync %.3f | joined %.3f | summed %.3f | factored %.3f | per device copy/update/done:", t_sync,
                t_join, t_sum, since());
        for (int d = 0; d < m->ndev; ++d) fprintf(stderr, " [%.3f %.3f %.3f asm %.3f]", t_copy[d], t_upd[d], t_done[d], m->t_asm[d]);
        fprintf(stderr, "\n");
    }
    return rcf;
}

}  // extern "C"
