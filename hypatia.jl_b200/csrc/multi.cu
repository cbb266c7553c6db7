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
#include <thread>

#include "ctx.cuh"

struct hyb_multi {
    int ndev = 0;
    std::vector<int> devices;
    hyb_ctx* root = nullptr;
    std::vector<hyb_ctx*> shard;         // shard[d] lives on devices[d]; shard[0] shares the root's G*Q2 rows
    std::vector<int64_t> row_lo, row_hi; // rows of s / z owned by shard d
    std::vector<double> t_asm;
    bool loaded = false;
};

namespace {

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
        if (devices[d] != devices[0]) {  // kernels on device d store into the root's memory
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[d], devices[0]);
            if (!can) {
                hyb_set_error("hyb_multi_create: device %d cannot access the memory of device %d (no P2P / NVLink path)", devices[d],
                              devices[0]);
                hyb_multi_destroy(m);
                return nullptr;
            }
            cudaSetDevice(devices[d]);
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                hyb_set_error("hyb_multi_create: cudaDeviceEnablePeerAccess(%d -> %d): %s", devices[d], devices[0],
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
    unsigned char handle[64];
    HYB_CHECK(hyb_peer_root_create(m->root, m->ndev, handle));  // ndev slots of ldS x N doubles on the root
    const ShardPlan plan = plan_shards(q, ncones, descs, m->ndev);
    const int64_t slot = m->root->ldS * N;
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
        s->peer_slot = m->root->peer_slots + (int64_t)d * slot;  // same process: a plain (peer) device pointer
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
    HYB_CUDA(cudaStreamSynchronize(r->st));
    std::vector<int> rc(m->ndev, 0);
    std::vector<std::string> err(m->ndev);
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
        int c = hyb_ensure_updated(s, s->loaded_scal);
        if (c == 0) c = hyb_schur_assemble_into(s, s->peer_slot, &m->t_asm[d]);
        if (c != 0) fail(c);
    };
    std::vector<std::thread> th;
    for (int d = 1; d < m->ndev; ++d) th.emplace_back(work, d);
    work(0);
    for (auto& t : th) t.join();
    for (int d = 0; d < m->ndev; ++d)
        if (rc[d] != 0) {
            hyb_set_error("device %d: %s", m->devices[d], err[d].c_str());
            return rc[d];
        }
    double ta = 0.0;
    for (int d = 0; d < m->ndev; ++d) ta = ta > m->t_asm[d] ? ta : m->t_asm[d];
    if (t_assemble_ms) *t_assemble_ms = ta;
    HYB_CHECK(hyb_peer_root_sum(r, t_reduce_ms));
    return hyb_schur_factor(r, t_fact_ms, fact_kind);
}

}  // extern "C"
