// Host side of libmgb200.so: handles, model packing, the constraint-program
// compiler, FITPACK-compatible B-spline basis, and the host-buffer entry point.
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>

#include <stddef.h>

#include "mgb_internal.cuh"

// ---------------------------------------------------------------------------
// errors / counters
// ---------------------------------------------------------------------------
static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void mgb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void mgb_count_launch(int n) { g_launches += n; }

extern "C" const char* mgb_last_error(void) { return g_err; }
extern "C" int64_t mgb_launch_count(void) { return g_launches.load(); }
extern "C" const char* mgb_version(void) { return "mgb200 0.1 (sm_100a)"; }

// ---------------------------------------------------------------------------
// B-spline basis: FITPACK splev / fpbspl semantics (cubic), including the
// clamp of the interval index that makes evaluation outside the knot range an
// extrapolation of the end polynomial piece (reference quirk Q1:
// motion_primitive.py:233 samples t = F, one past the last knot).
// ---------------------------------------------------------------------------
extern "C" int mgb_spline_basis(const double* knots, int32_t n_knots, double t, int32_t* span, double w[4]) {
    MGB_REQUIRE(knots && span && w, "null argument");
    const int k = 3;
    MGB_REQUIRE(n_knots >= 2 * (k + 1), "need at least %d knots", 2 * (k + 1));
    const int nk1 = n_knots - k - 1;
    int l = k;  // knots[l] <= t < knots[l+1], clamped to [k, nk1-1]
    while (l + 1 < nk1 && t >= knots[l + 1]) ++l;
    double h[4] = {1.0, 0.0, 0.0, 0.0}, hh[3];
    for (int j = 1; j <= k; ++j) {
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        for (int i = 0; i < j; ++i) {
            const int li = l + i + 1, lj = li - j;
            const double den = knots[li] - knots[lj];
            const double f = den != 0.0 ? hh[i] / den : 0.0;
            h[i] += f * (knots[li] - t);
            h[i + 1] = f * (t - knots[lj]);
        }
    }
    *span = l - k;
    for (int i = 0; i < 4; ++i) w[i] = h[i];
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// small fp64 helpers
// ---------------------------------------------------------------------------
static void quat_matrix_d(const double* q, double* R) {
    // transformations.quaternion_matrix
    const double n = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    if (n < 8.881784197001252e-16) {
        for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0;
        return;
    }
    const double s = 2.0 / n;
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    R[0] = 1.0 - s * (y * y + z * z); R[1] = s * (x * y - z * w);       R[2] = s * (x * z + y * w);
    R[3] = s * (x * y + z * w);       R[4] = 1.0 - s * (x * x + z * z); R[5] = s * (y * z - x * w);
    R[6] = s * (x * z - y * w);       R[7] = s * (y * z + x * w);       R[8] = 1.0 - s * (x * x + y * y);
}

static bool cholesky_lower(const double* A, int D, std::vector<double>& L) {
    L.assign((size_t)D * D, 0.0);
    for (int i = 0; i < D; ++i) {
        for (int j = 0; j <= i; ++j) {
            double s = A[(size_t)i * D + j];
            for (int k = 0; k < j; ++k) s -= L[(size_t)i * D + k] * L[(size_t)j * D + k];
            if (i == j) {
                if (s <= 0.0) return false;
                L[(size_t)i * D + i] = sqrt(s);
            } else {
                L[(size_t)i * D + j] = s / L[(size_t)j * D + j];
            }
        }
    }
    return true;
}

template <typename T>
static int upload(T** dst, const std::vector<T>& src) {
    *dst = nullptr;
    if (src.empty()) return MGB_OK;
    MGB_CUDA(cudaMalloc((void**)dst, sizeof(T) * src.size()));
    MGB_CUDA(cudaMemcpy(*dst, src.data(), sizeof(T) * src.size(), cudaMemcpyHostToDevice));
    return MGB_OK;
}

static void pack_pool_free(mgb_model* m);

// ---------------------------------------------------------------------------
// model
// ---------------------------------------------------------------------------
extern "C" int mgb_model_create(const mgb_model_desc* d, int device, mgb_model** out) {
    MGB_REQUIRE(d && out, "null argument");
    *out = nullptr;
    MGB_REQUIRE(d->n_frames > 0 && d->n_basis >= 4 && d->n_dim >= 7 && d->n_spatial > 0, "bad model shape");
    MGB_REQUIRE(d->n_latent >= d->n_spatial && d->n_gmm > 0, "bad GMM shape");
    MGB_REQUIRE((d->n_dim - 3) % 4 == 0, "n_dim must be 3 + 4 * joints");
    MGB_REQUIRE(d->eigen && d->mean && d->knots && d->translation_maxima && d->gmm_weights && d->gmm_means && d->gmm_covars,
                "null model array");
    MGB_REQUIRE(d->n_nodes > 0 && d->node_parent && d->node_type && d->node_channel && d->node_offset && d->node_rotation,
                "null skeleton array");
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev <= 0) {
        mgb_set_error("no CUDA device available (%s); libmgb200 has no CPU fallback", cudaGetErrorString(ce));
        return MGB_E_CUDA;
    }
    MGB_REQUIRE(device >= 0 && device < ndev, "device %d out of range (%d devices)", device, ndev);
    MGB_CUDA(cudaSetDevice(device));

    mgb_model* m = new mgb_model();
    m->device = device;
    m->F = d->n_frames; m->Nb = d->n_basis; m->P = d->n_dim; m->K = d->n_spatial; m->D = d->n_latent; m->G = d->n_gmm;
    m->M = m->Nb * m->P;
    const int K = m->K, M = m->M, P = m->P, D = m->D, G = m->G;
    m->E.resize((size_t)K * M);
    m->mean.resize(M);
    for (int j = 0; j < M; ++j) {
        const int pch = j % P;
        const double sc = pch < 3 ? d->translation_maxima[pch] : 1.0;  // motion_primitive.py:255
        m->mean[j] = d->mean[j] * sc;
        for (int k = 0; k < K; ++k) m->E[(size_t)k * M + j] = d->eigen[(size_t)k * M + j] * sc;
    }
    m->knots.assign(d->knots, d->knots + m->Nb + 4);
    // skeleton
    m->n_nodes = d->n_nodes;
    m->parent.assign(d->node_parent, d->node_parent + d->n_nodes);
    m->type.assign(d->node_type, d->node_type + d->n_nodes);
    m->channel.assign(d->node_channel, d->node_channel + d->n_nodes);
    m->offset.assign(d->node_offset, d->node_offset + 3 * d->n_nodes);
    m->fixed_rot.resize((size_t)9 * d->n_nodes);
    m->fixed_quat.resize((size_t)4 * d->n_nodes);
    m->children.assign(d->n_nodes, std::vector<int>());
    int n_root = 0;
    for (int i = 0; i < d->n_nodes; ++i) {
        const int pa = m->parent[i];
        if (pa < 0) {
            ++n_root;
        } else if (pa >= i) {
            mgb_set_error("skeleton nodes must be ordered parents first (node %d has parent %d)", i, pa);
            delete m;
            return MGB_E_INVALID;
        } else {
            m->children[pa].push_back(i);
        }
        if (m->channel[i] >= 0 && 3 + 4 * m->channel[i] + 4 > P) {
            mgb_set_error("node %d: channel %d outside the %d pose channels", i, m->channel[i], P);
            delete m;
            return MGB_E_INVALID;
        }
        quat_matrix_d(d->node_rotation + 4 * i, &m->fixed_rot[(size_t)9 * i]);
        {
            const double* q = d->node_rotation + 4 * i;
            const double nq = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
            const double inv = nq < 8.881784197001252e-16 ? 0.0 : 1.0 / sqrt(nq);
            m->fixed_quat[(size_t)4 * i + 0] = inv == 0.0 ? 1.0 : q[0] * inv;
            for (int c = 1; c < 4; ++c) m->fixed_quat[(size_t)4 * i + c] = inv == 0.0 ? 0.0 : q[c] * inv;
        }
    }
    if (n_root != 1 || m->parent[0] >= 0 || m->type[0] != MGB_NODE_ROOT) {
        mgb_set_error("skeleton must have exactly one root, at index 0, of type MGB_NODE_ROOT");
        delete m;
        return MGB_E_INVALID;
    }
    m->aligning_node = d->aligning_node;
    for (int i = 0; i < 3; ++i) m->aligning_dir[i] = d->aligning_dir[i];

    // GMM: precision Cholesky (sklearn _compute_precision_cholesky, 'full'), per-component constant
    m->NQ = (D + 3) / 4;
    m->DP = (D + 3) & ~3;
    const int DP = m->DP;
    std::vector<float4> gtab((size_t)G * m->NQ * (DP + 1), make_float4(0.f, 0.f, 0.f, 0.f));
    m->LS = (D + 3) & ~3;
    const int LS = m->LS;
    std::vector<float> gc(G), chol((size_t)G * D * LS, 0.f), means((size_t)G * D);
    std::vector<double> L, Pc((size_t)D * D);
    m->prec_chol.assign((size_t)G * D * D, 0.0);
    m->gmm_bias.assign((size_t)G * D, 0.0);
    m->chol64.assign((size_t)G * D * D, 0.0);
    m->means64.assign(d->gmm_means, d->gmm_means + (size_t)G * D);
    for (int g = 0; g < G; ++g) {
        if (!cholesky_lower(d->gmm_covars + (size_t)g * D * D, D, L)) {
            mgb_set_error("covariance of GMM component %d is not positive definite", g);
            delete m;
            return MGB_E_INVALID;
        }
        // X = L^-1 (lower), precision_chol = X^T (upper): y = (s - mu) . Pc
        std::fill(Pc.begin(), Pc.end(), 0.0);
        for (int c = 0; c < D; ++c) {  // solve L x = e_c
            for (int i = c; i < D; ++i) {
                double s = (i == c) ? 1.0 : 0.0;
                for (int k = c; k < i; ++k) s -= L[(size_t)i * D + k] * Pc[(size_t)c * D + k];  // Pc[c][k] = X[k][c]
                Pc[(size_t)c * D + i] = s / L[(size_t)i * D + i];                               // X[i][c] stored transposed
            }
        }
        double logdet = 0.0;
        for (int i = 0; i < D; ++i) logdet += log(Pc[(size_t)i * D + i]);
        gc[g] = (float)(log(d->gmm_weights[g]) + logdet - 0.5 * D * log(2.0 * M_PI));
        const double* mu = d->gmm_means + (size_t)g * D;
        for (int q = 0; q < m->NQ; ++q) {
            float4* row = &gtab[(size_t)(g * m->NQ + q) * (DP + 1)];
            double bias[4] = {0, 0, 0, 0};
            for (int k = 0; k < D; ++k) {
                float v[4] = {0, 0, 0, 0};
                for (int c = 0; c < 4; ++c) {
                    const int j = 4 * q + c;
                    if (j >= D) continue;
                    v[c] = (float)Pc[(size_t)k * D + j];
                    bias[c] -= mu[k] * Pc[(size_t)k * D + j];
                }
                row[1 + k] = make_float4(v[0], v[1], v[2], v[3]);
            }
            row[0] = make_float4((float)bias[0], (float)bias[1], (float)bias[2], (float)bias[3]);
        }
        for (int k = 0; k < D; ++k)
            for (int j = 0; j < D; ++j) {
                m->prec_chol[((size_t)g * D + k) * D + j] = Pc[(size_t)k * D + j];
                m->gmm_bias[(size_t)g * D + j] -= mu[k] * Pc[(size_t)k * D + j];
            }
        for (int i = 0; i < D; ++i)
            for (int j = 0; j <= i; ++j) {
                chol[((size_t)g * D + i) * LS + j] = (float)L[(size_t)i * D + j];
                m->chol64[((size_t)g * D + i) * D + j] = L[(size_t)i * D + j];
            }
        for (int i = 0; i < D; ++i) means[(size_t)g * D + i] = (float)mu[i];
    }
    m->gmm_tab_bytes = gtab.size() * sizeof(float4);

    std::vector<float> Ef(m->E.begin(), m->E.end()), meanf(m->mean.begin(), m->mean.end());
    std::vector<float> offf(m->offset.begin(), m->offset.end()), rotf(m->fixed_rot.begin(), m->fixed_rot.end());
    int rc = MGB_OK;
    if ((rc = upload(&m->d_E, Ef)) || (rc = upload(&m->d_mean, meanf)) || (rc = upload(&m->d_gmm_tab, gtab)) ||
        (rc = upload(&m->d_gmm_c, gc)) || (rc = upload(&m->d_chol, chol)) || (rc = upload(&m->d_means, means)) ||
        (rc = upload(&m->d_parent, m->parent)) || (rc = upload(&m->d_type, m->type)) || (rc = upload(&m->d_channel, m->channel)) ||
        (rc = upload(&m->d_offset, offf)) || (rc = upload(&m->d_fixed_rot, rotf))) {
        mgb_model_destroy(m);
        return rc;
    }
    // tensor-core back-projection operand (3xTF32 stage images); models too large for the
    // shared-memory plan simply keep the SIMT path
    if (mgb_backproject_tc_prepare(m) != MGB_OK) {
        cudaFree(m->d_Bsplit);
        m->d_Bsplit = nullptr;
    }
    if (mgb_score_tc_prepare_model(m) != MGB_OK) {
        cudaFree(m->d_Bgmm); cudaFree(m->d_Bsamp);
        m->d_Bgmm = nullptr; m->d_Bsamp = nullptr;
    }
    *out = m;
    return MGB_OK;
}

extern "C" void mgb_model_destroy(mgb_model* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    cudaFree(m->d_E); cudaFree(m->d_mean); cudaFree(m->d_gmm_tab); cudaFree(m->d_gmm_c); cudaFree(m->d_chol);
    cudaFree(m->d_means); cudaFree(m->d_parent); cudaFree(m->d_type); cudaFree(m->d_channel); cudaFree(m->d_offset);
    cudaFree(m->d_fixed_rot); cudaFree(m->d_Bsplit); cudaFree(m->d_Bgmm); cudaFree(m->d_Bsamp); cudaFree(m->d_ws);
    for (int b = 0; b < 2; ++b) {
        if (m->pinned[b]) cudaFreeHost(m->pinned[b]);
        cudaFree(m->d_stage[b]);
        if (m->ev_copied[b]) cudaEventDestroy(m->ev_copied[b]);
        if (m->ev_done[b]) cudaEventDestroy(m->ev_done[b]);
    }
    pack_pool_free(m);
    if (m->copy_stream) cudaStreamDestroy(m->copy_stream);
    if (m->compute_stream) cudaStreamDestroy(m->compute_stream);
    delete m;
}

// ---------------------------------------------------------------------------
// constraint-table blocks of a model (MgbPackBlock)
// ---------------------------------------------------------------------------
static int pack_block_acquire(const mgb_model* m, size_t bytes, MgbPackBlock** out) {
    std::lock_guard<std::mutex> lock(m->pool_mutex);
    if (!m->pack_stream) MGB_CUDA(cudaStreamCreateWithFlags(&m->pack_stream, cudaStreamNonBlocking));
    MgbPackBlock* best = nullptr;
    for (MgbPackBlock* b : m->pool)
        if (!b->busy && b->bytes >= bytes && (!best || b->bytes < best->bytes)) best = b;
    if (!best) {
        best = new MgbPackBlock();
        best->bytes = std::max<size_t>((bytes + 65535) & ~(size_t)65535, 65536);
        cudaError_t e = cudaMalloc((void**)&best->dev, best->bytes);
        if (e == cudaSuccess) e = cudaMallocHost((void**)&best->pin, best->bytes);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&best->copied, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&best->used, cudaEventDisableTiming);
        if (e != cudaSuccess) {
            mgb_set_error("constraint table block of %zu bytes: %s", best->bytes, cudaGetErrorString(e));
            cudaFree(best->dev);
            cudaFreeHost(best->pin);
            delete best;
            return MGB_E_CUDA;
        }
        m->pool.push_back(best);
    } else {
        MGB_CUDA(cudaEventSynchronize(best->copied));   // the pinned mirror is free again (long done in practice)
    }
    best->busy = true;
    *out = best;
    return MGB_OK;
}

static void pack_pool_free(mgb_model* m) {
    for (MgbPackBlock* b : m->pool) {
        cudaFree(b->dev);
        cudaFreeHost(b->pin);
        if (b->copied) cudaEventDestroy(b->copied);
        if (b->used) cudaEventDestroy(b->used);
        delete b;
    }
    m->pool.clear();
    if (m->pack_stream) cudaStreamDestroy(m->pack_stream);
    m->pack_stream = nullptr;
}

// ---------------------------------------------------------------------------
// constraint program compiler
// ---------------------------------------------------------------------------
namespace {

struct Action {
    int code;
    int a, b;
    std::vector<float> cst;
    int cst_at = -1;  // >= 0: the constants already sit in the pool at this offset (shared by several ops)
};

struct Compiler {
    const mgb_model* m;
    int n_times;
    std::vector<int> spans;
    std::vector<double> weights;  // [n_times + 1][4]; the extra slot is control point 0 (alignment)
    std::map<std::pair<int, int>, int> quad_of;  // (slot, group) -> quad; group -1 = translation, j = joint channel j
    std::vector<std::pair<int, int>> quads;
    std::vector<MgbOp> ops;
    std::vector<float> cst;
    std::vector<std::map<int, std::vector<Action>>> attach;  // [slot][node] -> actions
    std::vector<std::vector<Action>> free_actions;           // [slot] -> actions that read raw channels of the slot (no FK walk)
    int scratch_used = 0;
    int max_depth = 0;

    int quad(int slot, int group) {
        auto key = std::make_pair(slot, group);
        auto it = quad_of.find(key);
        if (it != quad_of.end()) return it->second;
        int q = (int)quads.size();
        quads.push_back(key);
        quad_of[key] = q;
        return q;
    }
    int push_cst(const std::vector<float>& v) {
        int off = (int)cst.size();
        cst.insert(cst.end(), v.begin(), v.end());
        return off;
    }
    void op(int code, int a, int b, const std::vector<float>& v) {
        MgbOp o;
        o.code = code; o.a = a; o.b = b; o.c = push_cst(v);
        ops.push_back(o);
    }
    void op_action(const Action& a) {
        if (a.cst_at < 0) { op(a.code, a.a, a.b, a.cst); return; }
        MgbOp o;
        o.code = a.code; o.a = a.a; o.b = a.b; o.c = a.cst_at;
        ops.push_back(o);
    }
    int alloc_scratch(int n) {
        int off = scratch_used;
        scratch_used += n;
        return off;
    }
    void add(int slot, int node, Action a) { attach[slot][node].push_back(std::move(a)); }

    bool subtree_needed(int node, const std::vector<char>& need) const { return need[node] != 0; }

    void emit_node(int slot, int node, const std::vector<char>& need, int depth) {
        const int type = m->type[node];
        const double* off = &m->offset[(size_t)3 * node];
        if (type == MGB_NODE_ROOT) {
            op(OP_ROOT, quad(slot, -1), quad(slot, m->channel[node] >= 0 ? m->channel[node] : 0),
               {(float)off[0], (float)off[1], (float)off[2]});
        } else if (type == MGB_NODE_JOINT && m->channel[node] >= 0) {
            op(OP_JOINT, quad(slot, m->channel[node]), 0, {(float)off[0], (float)off[1], (float)off[2]});
        } else {
            std::vector<float> v(7);
            bool ident = true;
            for (int i = 0; i < 4; ++i) {
                const double r = (type == MGB_NODE_END_SITE) ? (i == 0 ? 1.0 : 0.0) : m->fixed_quat[(size_t)4 * node + i];
                v[i] = (float)r;
                if (fabs(fabs(r) - (i == 0 ? 1.0 : 0.0)) > 0.0) ident = false;
            }
            for (int i = 0; i < 3; ++i) v[4 + i] = (float)off[i];
            op(OP_CONST, ident ? 1 : 0, 0, v);
        }
        auto it = attach[slot].find(node);
        if (it != attach[slot].end())
            for (const Action& a : it->second) op_action(a);
        std::vector<int> kids;
        for (int c : m->children[node])
            if (need[c]) kids.push_back(c);
        for (size_t i = 0; i < kids.size(); ++i) {
            const bool last = (i + 1 == kids.size());
            if (!last) {
                op(OP_PUSH, 0, 0, {});
                max_depth = std::max(max_depth, depth + 1);
            }
            emit_node(slot, kids[i], need, last ? depth : depth + 1);
            if (!last) op(OP_POP, 0, 0, {});
        }
    }

    void emit_slot(int slot) {
        // quads are numbered in emission order (the quads of a slot stay contiguous): the channel becomes a quad here
        for (const Action& a : free_actions[slot]) op(a.code, a.a, a.code == OP_JROT ? quad(slot, a.b) : a.b, a.cst);
        if (attach[slot].empty()) return;
        std::vector<char> need(m->n_nodes, 0);
        for (auto& kv : attach[slot]) {
            int nd = kv.first;
            while (nd >= 0 && !need[nd]) {
                need[nd] = 1;
                nd = m->parent[nd];
            }
        }
        emit_node(slot, 0, need, 0);
    }
};

}  // namespace

extern "C" int mgb_constraints_pack(const mgb_model* m, const mgb_constraint_desc* cd, int32_t n, const mgb_pose_point* points,
                                    int32_t n_points, const double* times, int32_t n_times, const int32_t* spans_in,
                                    const double* weights_in, const mgb_align_desc* align, mgb_constraints** out) {
    MGB_REQUIRE(m && out, "null argument");
    *out = nullptr;
    MGB_REQUIRE(n >= 0 && (n == 0 || cd), "bad constraint list");
    MGB_REQUIRE(n_times >= 0 && (n_times == 0 || times), "bad time list");
    MGB_REQUIRE((spans_in == nullptr) == (weights_in == nullptr), "spans and weights must be given together");
    Compiler C;
    C.m = m;
    C.n_times = n_times;
    C.spans.resize(n_times + 1);
    C.weights.resize((size_t)4 * (n_times + 1));
    for (int t = 0; t < n_times; ++t) {
        if (spans_in) {
            C.spans[t] = spans_in[t];
            for (int q = 0; q < 4; ++q) C.weights[4 * t + q] = weights_in[4 * t + q];
        } else {
            int rc = mgb_spline_basis(m->knots.data(), (int)m->knots.size(), times[t], &C.spans[t], &C.weights[4 * t]);
            if (rc != MGB_OK) return rc;
        }
        MGB_REQUIRE(C.spans[t] >= 0 && C.spans[t] + 4 <= m->Nb, "time %d: span %d out of range", t, C.spans[t]);
    }
    const int CP0 = n_times;  // pseudo slot: control point 0
    C.spans[CP0] = 0;
    C.weights[4 * CP0] = 1.0;
    C.weights[4 * CP0 + 1] = C.weights[4 * CP0 + 2] = C.weights[4 * CP0 + 3] = 0.0;
    C.attach.resize(n_times + 1);
    C.free_actions.resize(n_times + 1);

    auto check_node = [&](int nd) { return nd >= 0 && nd < m->n_nodes; };
    auto check_slot = [&](int s) { return s >= 0 && s < n_times; };
    std::map<int, int> lt_term_of_scratch;   // local-trajectory scratch offset -> constraint (residual-vector program)
    std::map<int, int> frame_term_of_scratch;   // trajectory / trajectory-set scratch offset -> constraint
    struct Final { int code, a, b; std::vector<float> cst; };
    std::vector<Final> finals;
    std::vector<MgbOp> begins;
    C.cst.push_back(0.f);  // offset 0 is a valid dummy for ops without constants

    for (int j = 0; j < n; ++j) {
        const mgb_constraint_desc& c = cd[j];
        const float w = (float)c.weight;
        switch (c.type) {
            case MGB_CT_POSITION:
            case MGB_CT_RELATIVE: {
                MGB_REQUIRE(check_node(c.node[0]) && check_slot(c.time_slot), "constraint %d: bad node / time slot", j);
                if (c.has_position) {
                    if (c.type == MGB_CT_POSITION) {
                        C.add(c.time_slot, c.node[0],
                              {OP_POS, j, (int)(c.axis_mask & 7u), {w, (float)c.target[0], (float)c.target[1], (float)c.target[2]}});
                    } else {
                        C.add(c.time_slot, c.node[0],
                              {OP_REL, j, 0, {w, (float)c.offset[0], (float)c.offset[1], (float)c.offset[2], (float)c.offset[3],
                                              (float)c.target[0], (float)c.target[1], (float)c.target[2]}});
                    }
                }
                if (c.has_orientation) {
                    // anim_utils quaternion_rotate_vector(q, [0,0,0,1]) for a general q, then unit length
                    const double qw = c.orient_q[0], qx = c.orient_q[1], qy = c.orient_q[2], qz = c.orient_q[3];
                    double v[3] = {2 * (qx * qz + qw * qy), 2 * (qy * qz - qw * qx), qw * qw - qx * qx - qy * qy + qz * qz};
                    double nv = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
                    MGB_REQUIRE(nv > 0, "constraint %d: zero orientation quaternion", j);
                    C.add(c.time_slot, c.node[0], {OP_ORIENT, j, 0, {w, (float)(v[0] / nv), (float)(v[1] / nv), (float)(v[2] / nv)}});
                }
                break;
            }
            case MGB_CT_DIR2D: {
                MGB_REQUIRE(check_slot(c.time_slot) && check_node(m->aligning_node), "constraint %d: bad time slot", j);
                const double len = sqrt(c.dir[0] * c.dir[0] + c.dir[1] * c.dir[1]);
                C.add(c.time_slot, m->aligning_node,
                      {OP_DIR2D, j, 0, {w, (float)m->aligning_dir[0], (float)m->aligning_dir[1], (float)m->aligning_dir[2],
                                        (float)c.dir[0], (float)c.dir[1], (float)len}});
                break;
            }
            case MGB_CT_LOOK_AT: {
                MGB_REQUIRE(check_node(c.node[0]) && check_slot(c.time_slot), "constraint %d: bad node / time slot", j);
                C.add(c.time_slot, c.node[0], {OP_LOOKAT, j, 0, {w, (float)c.target[0], (float)c.target[1], (float)c.target[2]}});
                break;
            }
            case MGB_CT_TWO_HANDS:
            case MGB_CT_FEET: {
                MGB_REQUIRE(check_node(c.node[0]) && check_node(c.node[1]) && check_slot(c.time_slot), "constraint %d: bad node / time slot", j);
                const int base = C.alloc_scratch(6);
                C.add(c.time_slot, c.node[0], {OP_STORE, base, 0, {}});
                C.add(c.time_slot, c.node[1], {OP_STORE, base + 3, 0, {}});
                if (c.type == MGB_CT_TWO_HANDS) {
                    finals.push_back({OP_TWOHANDS, j, base,
                                      {w, (float)c.center[0], (float)c.center[1], (float)c.center[2], (float)c.target[0], (float)c.target[1],
                                       (float)c.target[2], (float)c.target2[0], (float)c.target2[1], (float)c.target2[2]}});
                } else {
                    finals.push_back({OP_FEET, j, base,
                                      {w, (float)c.target[0], (float)c.target[1], (float)c.target[2], (float)c.target2[0], (float)c.target2[1],
                                       (float)c.target2[2], w}});
                }
                break;
            }
            case MGB_CT_POSE: {
                MGB_REQUIRE(check_slot(c.time_slot), "constraint %d: bad time slot", j);
                MGB_REQUIRE(c.n_points > 0 && c.n_points <= MGB_MAX_POSE_POINTS, "constraint %d: %d pose points (max %d)", j, c.n_points,
                            MGB_MAX_POSE_POINTS);
                MGB_REQUIRE(points && c.points_offset >= 0 && c.points_offset + c.n_points <= n_points, "constraint %d: pose points out of range", j);
                if (c.has_velocity) MGB_REQUIRE(check_slot(c.time_slot2), "constraint %d: bad second time slot", j);
                const int base = C.alloc_scratch(3 * c.n_points + 3);
                double W = 0, sax = 0, saz = 0;
                std::vector<float> v = {w, (float)c.n_points, c.has_velocity ? 1.f : 0.f, (float)c.velocity[0], (float)c.velocity[1],
                                        (float)c.velocity[2], 0.f, 0.f, 0.f};
                for (int q = 0; q < c.n_points; ++q) {
                    const mgb_pose_point& pp = points[c.points_offset + q];
                    MGB_REQUIRE(check_node(pp.node), "constraint %d: pose point %d has a bad node", j, q);
                    C.add(c.time_slot, pp.node, {OP_STORE, base + 3 * q, 0, {}});
                    W += pp.weight;
                    sax += pp.weight * pp.target[0];
                    saz += pp.weight * pp.target[2];
                    v.push_back((float)pp.target[0]);
                    v.push_back((float)pp.target[1]);
                    v.push_back((float)pp.target[2]);
                    v.push_back((float)pp.weight);
                }
                v[6] = (float)W; v[7] = (float)sax; v[8] = (float)saz;
                if (c.has_velocity) C.add(c.time_slot2, points[c.points_offset].node, {OP_STORE, base + 3 * c.n_points, 0, {}});
                finals.push_back({OP_POSE, j, base, v});
                break;
            }
            case MGB_CT_CA: {
                MGB_REQUIRE(check_node(c.node[0]) && c.n_slots > 0 && check_slot(c.time_slot) && check_slot(c.time_slot + c.n_slots - 1),
                            "constraint %d: bad node / slot range", j);
                const int sc = C.alloc_scratch(1);
                MgbOp b;
                b.code = OP_CA_BEGIN; b.a = sc; b.b = 0; b.c = 0;
                begins.push_back(b);
                for (int f = 0; f < c.n_slots; ++f)
                    C.add(c.time_slot + f, c.node[0],
                          {OP_CA_ACC, sc, (int)(c.axis_mask & 7u), {(float)c.target[0], (float)c.target[1], (float)c.target[2]}});
                finals.push_back({OP_CA_END, j, sc, {w}});
                break;
            }
            case MGB_CT_JOINT_ROT:
            case MGB_CT_QUAT_FRAME: {
                MGB_REQUIRE(check_slot(c.time_slot), "constraint %d: bad time slot", j);
                // target rotation matrix of a unit quaternion (transformations.quaternion_matrix), fp64
                auto rot_action = [&](int node, const double q[4], double inner) -> int {
                    MGB_REQUIRE(check_node(node) && m->channel[node] >= 0, "constraint %d: node %d has no quaternion channels", j, node);
                    const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
                    MGB_REQUIRE(n2 > 0, "constraint %d: zero target quaternion", j);
                    const double s2 = 2.0 / n2, qw = q[0], qx = q[1], qy = q[2], qz = q[3];
                    const double R[9] = {1 - (qy * qy + qz * qz) * s2, (qx * qy - qw * qz) * s2, (qx * qz + qw * qy) * s2,
                                         (qx * qy + qw * qz) * s2, 1 - (qx * qx + qz * qz) * s2, (qy * qz - qw * qx) * s2,
                                         (qx * qz - qw * qy) * s2, (qy * qz + qw * qx) * s2, 1 - (qx * qx + qy * qy) * s2};
                    std::vector<float> v = {w, (float)inner};
                    for (int i = 0; i < 9; ++i) v.push_back((float)R[i]);
                    v.push_back(m->type[node] == MGB_NODE_ROOT ? 1.f : 0.f);
                    C.free_actions[c.time_slot].push_back({OP_JROT, j, m->channel[node], v});   // b: channel now, quad at emission
                    return MGB_OK;
                };
                if (c.type == MGB_CT_JOINT_ROT) {
                    if (int rc = rot_action(c.node[0], c.orient_q, 1.0)) return rc;
                } else {
                    MGB_REQUIRE(c.n_points > 0 && points && c.points_offset >= 0 && c.points_offset + c.n_points <= n_points,
                                "constraint %d: pose points out of range", j);
                    for (int q = 0; q < c.n_points; ++q) {
                        const mgb_pose_point& pp = points[c.points_offset + q];
                        const double v2 = pp.target[0] * pp.target[0] + pp.target[1] * pp.target[1] + pp.target[2] * pp.target[2];
                        MGB_REQUIRE(v2 <= 1.0 + 1e-9, "constraint %d: pose point %d is not the vector part of a unit quaternion", j, q);
                        const double quat[4] = {sqrt(std::max(0.0, 1.0 - v2)), pp.target[0], pp.target[1], pp.target[2]};
                        if (int rc = rot_action(pp.node, quat, pp.weight)) return rc;
                    }
                }
                break;
            }
            case MGB_CT_LOCAL_TRAJ: {
                MGB_REQUIRE(check_node(c.node[0]) && c.n_slots > 0 && check_slot(c.time_slot) && check_slot(c.time_slot + c.n_slots - 1),
                            "constraint %d: bad node / slot range", j);
                MGB_REQUIRE(c.n_points >= 3 && points && c.points_offset >= 0 && c.points_offset + c.n_points <= n_points && c.offset[1] > 0,
                            "constraint %d: bad trajectory table", j);
                const int sc = C.alloc_scratch(6);
                MgbOp b;
                b.code = OP_LT_BEGIN; b.a = sc; b.b = 0; b.c = C.push_cst({(float)c.offset[0]});
                begins.push_back(b);
                const int nt = c.n_points - 2;
                std::vector<float> v = {(float)(1.0 / c.offset[1]), (float)nt, (float)(nt * c.offset[1]), w};
                lt_term_of_scratch[sc] = j;
                for (int q = 0; q < c.n_points; ++q) {
                    v.push_back((float)points[c.points_offset + q].target[0]);
                    v.push_back((float)points[c.points_offset + q].target[2]);
                }
                // one copy of the table in the constant pool, shared by the per-frame ops
                const int tab_off = C.push_cst(v);
                for (int f = 0; f < c.n_slots; ++f) {
                    Action a{OP_LT_ACC, sc, 0, {}};
                    a.cst_at = tab_off;
                    C.add(c.time_slot + f, c.node[0], a);
                }
                finals.push_back({OP_LT_END, j, sc, {w}});
                break;
            }
            case MGB_CT_TRAJ: {
                MGB_REQUIRE(check_node(c.node[0]) && c.n_slots > 0 && check_slot(c.time_slot) && check_slot(c.time_slot + c.n_slots - 1),
                            "constraint %d: bad node / slot range", j);
                MGB_REQUIRE(c.n_points >= 2 && points && c.points_offset >= 0 && c.points_offset + c.n_points <= n_points,
                            "constraint %d: bad trajectory table", j);
                MGB_REQUIRE(c.offset[0] >= 0.0 && c.offset[0] <= 1.0, "constraint %d: the start parameter must lie in [0, 1]", j);
                const int sc = C.alloc_scratch(2);
                MgbOp b;
                b.code = OP_SCR_INIT; b.a = sc; b.b = 2; b.c = C.push_cst({0.f, (float)c.offset[0]});
                begins.push_back(b);
                const int nt = c.n_points - 1;
                std::vector<float> v = {(float)nt, w, 0.f, 0.f};
                for (int q = 0; q < c.n_points; ++q)
                    for (int i = 0; i < 3; ++i) v.push_back((float)points[c.points_offset + q].target[i]);
                const int tab_off = C.push_cst(v);
                frame_term_of_scratch[sc] = j;
                for (int f = 0; f < c.n_slots; ++f) {
                    Action a{OP_TR_ACC, sc, 0, {}};
                    a.cst_at = tab_off;
                    C.add(c.time_slot + f, c.node[0], a);
                }
                finals.push_back({OP_MEAN_END, j, sc, {w, (float)(1.0 / c.n_slots)}});
                break;
            }
            case MGB_CT_TRAJ_SET: {
                const int J = c.node[1];
                MGB_REQUIRE(J >= 1 && J <= 8, "constraint %d: a trajectory set holds 1 .. 8 joint trajectories", j);
                MGB_REQUIRE(c.n_slots > 0 && check_slot(c.time_slot) && check_slot(c.time_slot + c.n_slots - 1), "constraint %d: bad slot range", j);
                MGB_REQUIRE(points && c.points_offset >= 0 && c.points_offset + c.n_points <= n_points, "constraint %d: bad trajectory tables", j);
                const int sc = C.alloc_scratch(3 + 5 * J);
                std::vector<float> init(3 + 5 * J, 0.f);
                int at = c.points_offset;
                const int end = c.points_offset + c.n_points;
                std::vector<int> joint_node(J), tab_off(J);
                for (int jj = 0; jj < J; ++jj) {
                    MGB_REQUIRE(at + 2 <= end, "constraint %d: trajectory %d: header missing", j, jj);
                    const mgb_pose_point& h0 = points[at];
                    const mgb_pose_point& h1 = points[at + 1];
                    const double step = h0.target[0], arc0 = h0.target[1];
                    const int nt = (int)h0.target[2];
                    MGB_REQUIRE(check_node(h0.node) && step > 0.0 && nt >= 1 && at + 2 + nt + 2 <= end, "constraint %d: trajectory %d: bad header / table", j, jj);
                    joint_node[jj] = h0.node;
                    init[3 + 5 * jj] = (float)arc0;
                    const bool ranged = h1.target[2] != 0.0;
                    // no range set: never active (trajectory_constraint.py:150-151)
                    std::vector<float> v = {(float)(1.0 / step), (float)nt, (float)(nt * step), ranged ? (float)h1.target[0] : 1.f,
                                            ranged ? (float)h1.target[1] : 0.f, (float)(1.0 / (3.0 * J)), w, 0.f};
                    for (int q = 0; q < nt + 2; ++q)
                        for (int i = 0; i < 3; ++i) v.push_back((float)points[at + 2 + q].target[i]);
                    tab_off[jj] = C.push_cst(v);
                    at += 2 + nt + 2;
                }
                MgbOp b;
                b.code = OP_SCR_INIT; b.a = sc; b.b = 3 + 5 * J; b.c = C.push_cst(init);
                begins.push_back(b);
                frame_term_of_scratch[sc] = j;
                for (int f = 0; f < c.n_slots; ++f)
                    for (int jj = 0; jj < J; ++jj) {
                        Action a{OP_TS_ACC, sc, jj << 1, {}};
                        a.cst_at = tab_off[jj];
                        C.add(c.time_slot + f, joint_node[jj], a);
                    }
                finals.push_back({OP_MEAN_END, j, sc, {w, (float)(1.0 / c.n_slots)}});
                break;
            }
            default:
                mgb_set_error("constraint %d: type %d is outside the device path", j, c.type);
                return MGB_E_UNSUPPORTED;
        }
    }
    MGB_REQUIRE(C.scratch_used <= MGB_SCRATCH_FLOATS, "constraint set needs %d scratch floats (max %d)", C.scratch_used, MGB_SCRATCH_FLOATS);

    // ---- program: [CA begins] [alignment prelude] [slots ...] [finals] -------
    for (const MgbOp& b : begins) C.ops.push_back(b);
    const bool do_align = align && align->mode == 1;
    if (do_align) {
        MGB_REQUIRE(check_node(m->aligning_node), "bad aligning node");
        const double PI = M_PI;
        C.add(CP0, m->aligning_node,
              {OP_ALIGN, C.quad(CP0, -1), 0,
               {(float)m->aligning_dir[0], (float)m->aligning_dir[1], (float)m->aligning_dir[2], (float)(align->prev_heading_deg * PI / 180.0),
                (float)align->prev_root[0], (float)align->prev_root[2]}});
        C.emit_slot(CP0);
    }
    for (int t = 0; t < n_times; ++t) C.emit_slot(t);
    for (const Final& f : finals) C.op(f.code, f.a, f.b, f.cst);
    MGB_REQUIRE(C.max_depth <= MGB_STACK_DEPTH, "skeleton walk needs a transform stack of %d (max %d)", C.max_depth, MGB_STACK_DEPTH);

    // ---- contracted quad table (fp64 accumulate, fp32 store) ------------------
    const int K = m->K, P = m->P, M = m->M;
    mgb_constraints* c = new mgb_constraints();
    c->model = m;
    c->n_terms = n;
    c->n_times = n_times;
    c->n_quads = (int)C.quads.size();
    const int KP = (K + 3) & ~3;  // rows padded with zeros to a multiple of 4; entry 0 of a quad is the bias (mean) row
    c->KP = KP;
    c->tab.assign((size_t)c->n_quads * (KP + 1) * 4, 0.f);
    c->tab64.assign((size_t)c->n_quads * (KP + 1) * 4, 0.0);
    // row k of the basis outermost: E is [K, M] row-major with M = n_coeffs * P (45 KB per row at L), so the gather over the
    // quads' columns stays inside one row at a time (this runs once per motion-primitive step, on the planner's thread)
    for (int k = 0; k <= K; ++k) {
        const double* basis = (k < K) ? m->E.data() + (size_t)k * M : m->mean.data();
        const int row = (k < K) ? 1 + k : 0;
        for (int q = 0; q < c->n_quads; ++q) {
            const int slot = C.quads[q].first, group = C.quads[q].second;
            const int ch0 = group < 0 ? 0 : 3 + 4 * group;
            const int nch = group < 0 ? 3 : 4;
            for (int ch = 0; ch < nch; ++ch) {
                double acc = 0.0;
                for (int i = 0; i < 4; ++i) acc += C.weights[4 * slot + i] * basis[(C.spans[slot] + i) * P + ch0 + ch];
                c->tab[((size_t)q * (KP + 1) + row) * 4 + ch] = (float)acc;
                c->tab64[((size_t)q * (KP + 1) + row) * 4 + ch] = acc;
            }
        }
    }
    {
        // trajectory sets: the J joint ops of a frame arrive in the order of the skeleton walk (the same in every frame); the
        // J-th one closes the frame
        std::map<int, int> met;
        for (MgbOp& o : C.ops)
            if (o.code == OP_TS_ACC) {
                const int J = cd[frame_term_of_scratch[o.a]].node[1];
                if (++met[o.a] % J == 0) o.b |= 1;
            }
    }
    c->ops_fused = C.ops;
    // ---- residual-vector program: residual layout = constraints in list order, each with the residuals
    //      its class returns from get_residual_vector_spline (1; two hands 3; pose / quaternion frame one per point; local
    //      trajectory one per frame); fields that named a term now name a residual ---------------------------------------
    {
        std::vector<int> res_base(n + 1, 0);
        for (int j = 0; j < n; ++j) {
            int cnt = 1;
            if (cd[j].type == MGB_CT_TWO_HANDS) cnt = 3;
            else if (cd[j].type == MGB_CT_POSE || cd[j].type == MGB_CT_QUAT_FRAME) cnt = cd[j].n_points;
            else if (cd[j].type == MGB_CT_LOCAL_TRAJ || cd[j].type == MGB_CT_TRAJ || cd[j].type == MGB_CT_TRAJ_SET) cnt = cd[j].n_slots;
            res_base[j + 1] = res_base[j] + cnt;
        }
        c->n_resid = res_base[n];
        c->ops_resid = C.ops;
        std::vector<int> seen(n, 0);   // JROT ops / local-trajectory frames of a constraint met so far (program order = point / frame order)
        for (MgbOp& o : c->ops_resid) {
            switch (o.code) {
                case OP_POS: case OP_ORIENT: case OP_REL: case OP_DIR2D: case OP_LOOKAT: case OP_FEET: case OP_CA_END:
                case OP_TWOHANDS: case OP_POSE:
                    o.a = res_base[o.a];
                    break;
                case OP_JROT:
                    o.a = res_base[o.a] + (cd[o.a].type == MGB_CT_QUAT_FRAME ? seen[o.a]++ : 0);
                    break;
                case OP_LT_ACC: {
                    const int term = lt_term_of_scratch[o.a];
                    o.b = res_base[term] + seen[term]++;
                    break;
                }
                case OP_TR_ACC: {
                    const int term = frame_term_of_scratch[o.a];
                    o.b = res_base[term] + seen[term]++;
                    break;
                }
                case OP_TS_ACC: {
                    const int term = frame_term_of_scratch[o.a];
                    o.b = (o.b & 255) | ((res_base[term] + seen[term]) << 8);
                    if (o.b & 1) ++seen[term];
                    break;
                }
                default:
                    break;
            }
        }
    }
    c->quad_slot.resize(C.quads.size());
    for (size_t q = 0; q < C.quads.size(); ++q) c->quad_slot[q] = C.quads[q].first;
    // frames-mode program: quads index channels of the materialised frames
    int cp0_slot = -1;
    for (int t = 0; t < n_times; ++t)
        if (C.spans[t] == 0 && C.weights[4 * t] == 1.0 && C.weights[4 * t + 1] == 0.0 && C.weights[4 * t + 2] == 0.0 && C.weights[4 * t + 3] == 0.0)
            cp0_slot = t;
    bool frames_ok = true;
    c->ops_frames = C.ops;
    for (MgbOp& o : c->ops_frames) {
        auto remap = [&](int q) -> int {
            int slot = C.quads[q].first, group = C.quads[q].second;
            if (slot == CP0) {
                if (cp0_slot < 0) { frames_ok = false; return 0; }
                slot = cp0_slot;
            }
            return slot * P + (group < 0 ? 0 : 3 + 4 * group);
        };
        if (o.code == OP_ROOT) { o.a = remap(o.a); o.b = remap(o.b); }
        else if (o.code == OP_JOINT || o.code == OP_ALIGN) { o.a = remap(o.a); }
        else if (o.code == OP_JROT) { o.b = remap(o.b); }
    }
    if (!frames_ok) c->ops_frames.clear();
    c->cst = C.cst;
    c->n_ops = (int)C.ops.size();
    c->n_cst = (int)C.cst.size();
    c->times.assign(times, times + n_times);
    c->spans.assign(C.spans.begin(), C.spans.begin() + n_times);
    c->weights.assign(C.weights.begin(), C.weights.begin() + (size_t)4 * n_times);

    MGB_CUDA(cudaSetDevice(m->device));
    // ---- one block, one asynchronous copy: [tab4 | ops_fused | ops_frames | cst], each part 256-byte aligned ----------
    auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t sz_tab = (size_t)c->n_quads * (KP + 1) * sizeof(float4), sz_of = c->ops_fused.size() * sizeof(MgbOp),
                 sz_ofr = c->ops_frames.size() * sizeof(MgbOp), sz_cst = c->cst.size() * sizeof(float), sz_ors = c->ops_resid.size() * sizeof(MgbOp);
    const size_t o_tab = 0, o_of = al(sz_tab), o_ofr = o_of + al(sz_of), o_cst = o_ofr + al(sz_ofr), o_ors = o_cst + al(sz_cst),
                 total = o_ors + al(sz_ors);
    MgbPackBlock* blk = nullptr;
    if (int rc = pack_block_acquire(m, total, &blk)) {
        delete c;
        return rc;
    }
    c->block = blk;
    {
        float* t = reinterpret_cast<float*>(blk->pin + o_tab);
        memcpy(t, c->tab.data(), sz_tab);    // tab is [n_quads][KP + 1][4] floats = the float4 table
        if (sz_of) memcpy(blk->pin + o_of, c->ops_fused.data(), sz_of);
        if (sz_ofr) memcpy(blk->pin + o_ofr, c->ops_frames.data(), sz_ofr);
        if (sz_cst) memcpy(blk->pin + o_cst, c->cst.data(), sz_cst);
        if (sz_ors) memcpy(blk->pin + o_ors, c->ops_resid.data(), sz_ors);
    }
    c->d_tab = sz_tab ? reinterpret_cast<float4*>(blk->dev + o_tab) : nullptr;
    c->d_ops_fused = sz_of ? reinterpret_cast<MgbOp*>(blk->dev + o_of) : nullptr;
    c->d_ops_frames = sz_ofr ? reinterpret_cast<MgbOp*>(blk->dev + o_ofr) : nullptr;
    c->d_cst = sz_cst ? reinterpret_cast<float*>(blk->dev + o_cst) : nullptr;
    c->d_ops_resid = sz_ors ? reinterpret_cast<MgbOp*>(blk->dev + o_ors) : nullptr;
    {
        cudaError_t e = cudaStreamWaitEvent(m->pack_stream, blk->used, 0);      // launches that still read the block's previous tenant
        if (e == cudaSuccess) e = cudaMemcpyAsync(blk->dev, blk->pin, total, cudaMemcpyHostToDevice, m->pack_stream);
        if (e == cudaSuccess) e = cudaEventRecord(blk->copied, m->pack_stream);
        if (e != cudaSuccess) {
            mgb_set_error("constraint table upload failed: %s", cudaGetErrorString(e));
            mgb_constraints_destroy(c);
            return MGB_E_CUDA;
        }
    }
    if (mgb_score_tc_prepare_constraints(m, c) != MGB_OK) {  // e.g. a slot with more quads than one chunk: SIMT path only
        cudaFree(c->d_Bq); cudaFree(c->d_ops_tc); cudaFree(c->d_phase_tc);
        c->d_Bq = nullptr; c->d_ops_tc = nullptr; c->d_phase_tc = nullptr;
    }
    *out = c;
    return MGB_OK;
}

extern "C" void mgb_constraints_destroy(mgb_constraints* c) {
    if (!c) return;
    cudaFree(c->d_Bq); cudaFree(c->d_ops_tc); cudaFree(c->d_phase_tc);
    if (c->block && c->model) {     // back to the model's pool; kernels in flight are covered by the block's `used` event
        std::lock_guard<std::mutex> lock(c->model->pool_mutex);
        c->block->busy = false;
    }
    delete c;
}

int mgb_constraints_before_launch(const mgb_constraints* c, cudaStream_t stream) {
    if (c && c->block) MGB_CUDA(cudaStreamWaitEvent(stream, c->block->copied, 0));
    return MGB_OK;
}
int mgb_constraints_after_launch(const mgb_constraints* c, cudaStream_t stream) {
    if (c && c->block) MGB_CUDA(cudaEventRecord(c->block->used, stream));
    return MGB_OK;
}

extern "C" int32_t mgb_constraints_n_terms(const mgb_constraints* c) { return c ? c->n_terms : 0; }

extern "C" int mgb_constraints_info(const mgb_constraints* c, int32_t* n_quads, int32_t* n_ops, int32_t* n_times) {
    MGB_REQUIRE(c, "null handle");
    if (n_quads) *n_quads = c->n_quads;
    if (n_ops) *n_ops = c->n_ops;
    if (n_times) *n_times = c->n_times;
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// scoring entry points
// ---------------------------------------------------------------------------
extern "C" int mgb_score(const mgb_model* m, const mgb_constraints* c, const void* d_s, int32_t s_is_f64, int64_t n, int64_t index_base,
                         float err_scale, float quality_scale, float magnitude_scale, float* d_cost, float* d_terms, int32_t k,
                         float* d_best_cost, int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream) {
    MGB_REQUIRE(c, "null constraints handle");
    return mgb_launch_score(m, c, 0, d_s, s_is_f64, nullptr, nullptr, 0, n, index_base, err_scale, quality_scale, magnitude_scale, d_cost,
                            d_terms, k, d_best_cost, d_best_idx, d_workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int mgb_sample_score(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base, const int64_t* h_counts,
                                uint64_t seed, float err_scale, float quality_scale, float magnitude_scale, float* d_cost, int32_t k,
                                float* d_best_cost, int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream) {
    MGB_REQUIRE(c, "null constraints handle");
    return mgb_launch_score(m, c, 2, nullptr, 0, nullptr, h_counts, seed, n, index_base, err_scale, quality_scale, magnitude_scale, d_cost,
                            nullptr, k, d_best_cost, d_best_idx, d_workspace, workspace_bytes, (cudaStream_t)stream);
}

static int exchange_from_desc(const mgb_exchange_desc* x, MgbExchange& xch) {
    memset(&xch, 0, sizeof(xch));
    if (!x) return MGB_OK;
    MGB_REQUIRE(x->d_peer_slots && x->d_err, "exchange: null argument");
    MGB_REQUIRE(x->world >= 1 && x->world <= 32 && x->rank >= 0 && x->rank < x->world, "exchange: rank / world out of range (one box: world <= 32)");
    MGB_REQUIRE(x->seq != 0, "exchange: seq 0 is the cleared state of the slots; start at 1");
    xch.peer_slots = x->d_peer_slots;
    xch.rank = x->rank;
    xch.world = x->world;
    xch.seq = x->seq;
    xch.spin_limit = x->spin_limit > 0 ? x->spin_limit : 20000000;
    xch.err = x->d_err;
    return MGB_OK;
}

extern "C" int mgb_sample_score_sharded(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base, const int64_t* h_counts,
                                        uint64_t seed, float err_scale, float quality_scale, float magnitude_scale,
                                        const mgb_exchange_desc* x, float* d_best_cost, int64_t* d_best_idx, void* d_workspace,
                                        size_t workspace_bytes, void* stream) {
    MGB_REQUIRE(c, "null constraints handle");
    MgbExchange xch;
    if (int rc = exchange_from_desc(x, xch)) return rc;
    return mgb_launch_score(m, c, 2, nullptr, 0, nullptr, h_counts, seed, n, index_base, err_scale, quality_scale, magnitude_scale, nullptr,
                            nullptr, 1, d_best_cost, d_best_idx, d_workspace, workspace_bytes, (cudaStream_t)stream, nullptr, x ? &xch : nullptr);
}

// ---------------------------------------------------------------------------
// planner-facing entries: results (cost, index, winning row) come back to HOST memory in one copy
// ---------------------------------------------------------------------------
namespace {
constexpr int SCR_MAX_JOBS = 64;
constexpr int SCR_ROW = 256;    // floats per winning row (latent dimension <= 256)
struct ScrResult {
    float cost;
    int pad;
    long long idx;
    float row[SCR_ROW];
};
struct DevScratch {
    std::mutex mu;
    bool ready = false;
    unsigned char* d_ws = nullptr;
    size_t ws_stride = 0;
    ScrResult* d_res = nullptr;
    ScrResult* h_res = nullptr;          // pinned
    ScoreParams* d_params = nullptr;
    ScoreParams* h_params = nullptr;     // pinned
    int* d_first = nullptr;
    int* h_first = nullptr;              // pinned
};
DevScratch g_scratch[MGB_MAX_DEVICES];

int scratch_for(int device, DevScratch** out) {
    MGB_REQUIRE(device >= 0 && device < MGB_MAX_DEVICES, "device %d out of range", device);
    DevScratch& s = g_scratch[device];
    if (!s.ready) {
        MGB_CUDA(cudaSetDevice(device));
        s.ws_stride = (mgb_score_workspace_bytes(0, 1) + 255) & ~(size_t)255;
        MGB_CUDA(cudaMalloc((void**)&s.d_ws, s.ws_stride * SCR_MAX_JOBS));
        MGB_CUDA(cudaMemset(s.d_ws, 0, s.ws_stride * SCR_MAX_JOBS));
        MGB_CUDA(cudaMalloc((void**)&s.d_res, sizeof(ScrResult) * SCR_MAX_JOBS));
        MGB_CUDA(cudaMallocHost((void**)&s.h_res, sizeof(ScrResult) * SCR_MAX_JOBS));
        MGB_CUDA(cudaMalloc((void**)&s.d_params, sizeof(ScoreParams) * SCR_MAX_JOBS));
        MGB_CUDA(cudaMallocHost((void**)&s.h_params, sizeof(ScoreParams) * SCR_MAX_JOBS));
        MGB_CUDA(cudaMalloc((void**)&s.d_first, sizeof(int) * (SCR_MAX_JOBS + 1)));
        MGB_CUDA(cudaMallocHost((void**)&s.h_first, sizeof(int) * (SCR_MAX_JOBS + 1)));
        s.ready = true;
    }
    *out = &s;
    return MGB_OK;
}
}  // namespace

extern "C" int mgb_sample_score_best(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base, const int64_t* h_counts,
                                     uint64_t seed, float err_scale, float quality_scale, float magnitude_scale, float* h_best_cost,
                                     int64_t* h_best_idx, float* h_best_row, void* stream_) {
    return mgb_sample_score_best_sharded(m, c, n, index_base, h_counts, seed, err_scale, quality_scale, magnitude_scale, nullptr, h_best_cost,
                                         h_best_idx, h_best_row, stream_);
}

extern "C" int mgb_sample_score_best_sharded(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base, const int64_t* h_counts,
                                             uint64_t seed, float err_scale, float quality_scale, float magnitude_scale,
                                             const mgb_exchange_desc* x, float* h_best_cost, int64_t* h_best_idx, float* h_best_row,
                                             void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(m && h_best_cost && h_best_idx, "null argument");
    MgbExchange xch;
    if (int rcx = exchange_from_desc(x, xch)) return rcx;
    MGB_REQUIRE(!h_best_row || m->D <= SCR_ROW, "latent dimension %d exceeds the result record (%d)", m->D, SCR_ROW);
    DevScratch* s = nullptr;
    if (int rc = scratch_for(m->device, &s)) return rc;
    std::lock_guard<std::mutex> lock(s->mu);
    MGB_CUDA(cudaSetDevice(m->device));
    int rc = mgb_launch_score(m, c, 2, nullptr, 0, nullptr, h_counts, seed, n, index_base, err_scale, quality_scale, magnitude_scale, nullptr,
                              nullptr, 1, &s->d_res[0].cost, (int64_t*)&s->d_res[0].idx, s->d_ws, s->ws_stride, stream,
                              h_best_row ? s->d_res[0].row : nullptr, x ? &xch : nullptr);
    if (rc != MGB_OK) return rc;
    const size_t bytes = offsetof(ScrResult, row) + (h_best_row ? sizeof(float) * m->D : 0);
    MGB_CUDA(cudaMemcpyAsync(s->h_res, s->d_res, bytes, cudaMemcpyDeviceToHost, stream));
    MGB_CUDA(cudaStreamSynchronize(stream));
    *h_best_cost = s->h_res[0].cost;
    *h_best_idx = s->h_res[0].idx;
    if (h_best_row) memcpy(h_best_row, s->h_res[0].row, sizeof(float) * m->D);
    return MGB_OK;
}

extern "C" int mgb_sample_score_multi(const mgb_option* options, int32_t n_options, float* h_best_cost, int64_t* h_best_idx,
                                      float* h_best_rows, int32_t row_stride, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(options && n_options > 0 && h_best_cost && h_best_idx, "null argument");
    MGB_REQUIRE(n_options <= SCR_MAX_JOBS, "at most %d options per launch", SCR_MAX_JOBS);
    const int device = options[0].model ? options[0].model->device : -1;
    DevScratch* s = nullptr;
    if (int rc = scratch_for(device, &s)) return rc;
    std::lock_guard<std::mutex> lock(s->mu);
    MGB_CUDA(cudaSetDevice(device));
    std::vector<MgbJob> jobs(n_options);
    for (int j = 0; j < n_options; ++j) {
        const mgb_option& o = options[j];
        MGB_REQUIRE(o.model && o.counts, "option %d: null model / counts", j);
        MGB_REQUIRE(!h_best_rows || (o.model->D <= SCR_ROW && o.model->D <= row_stride), "option %d: latent dimension %d exceeds the row stride", j,
                    o.model->D);
        jobs[j] = {o.model, o.constraints, o.n, o.index_base, o.counts, o.seed, o.err_scale, o.quality_scale, o.magnitude_scale,
                   &s->d_res[j].cost, (int64_t*)&s->d_res[j].idx, h_best_rows ? s->d_res[j].row : nullptr};
    }
    int rc = mgb_launch_score_multi(jobs.data(), n_options, s->h_params, s->h_first, s->d_params, s->d_first, s->d_ws, s->ws_stride, stream);
    if (rc != MGB_OK) return rc;
    MGB_CUDA(cudaMemcpyAsync(s->h_res, s->d_res, sizeof(ScrResult) * n_options, cudaMemcpyDeviceToHost, stream));
    MGB_CUDA(cudaStreamSynchronize(stream));
    for (int j = 0; j < n_options; ++j) {
        h_best_cost[j] = s->h_res[j].cost;
        h_best_idx[j] = s->h_res[j].idx;
        if (h_best_rows) memcpy(h_best_rows + (size_t)j * row_stride, s->h_res[j].row, sizeof(float) * options[j].model->D);
    }
    return MGB_OK;
}

extern "C" int mgb_node_costs(const mgb_model* m, const mgb_constraints* c, const void* d_points, int32_t is_f64, int64_t n,
                              float err_scale, float quality_scale, float magnitude_scale, float* d_cost, void* stream) {
    MGB_REQUIRE(c, "null constraints handle");
    MGB_REQUIRE(n == 0 || (d_points && d_cost), "null argument");
    // the tree's point table scored like any candidate batch, no selection: the walk needs every cost
    return mgb_launch_score(m, c, 0, d_points, is_f64, nullptr, nullptr, 0, n, 0, err_scale, quality_scale, magnitude_scale, d_cost, nullptr,
                            0, nullptr, nullptr, nullptr, 0, (cudaStream_t)stream);
}

extern "C" int mgb_logpdf(const mgb_model* m, const float* d_s, int64_t n, float* d_logpdf, void* stream) {
    MGB_REQUIRE(m && (n == 0 || (d_s && d_logpdf)), "null argument");
    // the score kernel with an empty program: cost = quality_scale * (-log p(s)) with quality_scale = -1
    return mgb_launch_score(m, nullptr, 0, d_s, 0, nullptr, nullptr, 0, n, 0, 0.f, -1.f, 0.f, d_logpdf, nullptr, 0, nullptr, nullptr, nullptr, 0,
                            (cudaStream_t)stream);
}

extern "C" int mgb_score_frames(const mgb_model* m, const mgb_constraints* c, const float* d_frames, const float* d_s, int64_t n,
                                int64_t index_base, float err_scale, float quality_scale, float magnitude_scale, float* d_cost,
                                float* d_terms, int32_t k, float* d_best_cost, int64_t* d_best_idx, void* d_workspace,
                                size_t workspace_bytes, void* stream) {
    MGB_REQUIRE(d_frames, "d_frames is NULL");
    MGB_REQUIRE(c && (c->n_ops == 0 || !c->ops_frames.empty()),
                "this constraint set needs control point 0 (alignment): include time 0 in the packed times to score from frames");
    return mgb_launch_score(m, c, 1, d_s, 0, d_frames, nullptr, 0, n, index_base, err_scale, quality_scale, magnitude_scale, d_cost, d_terms,
                            k, d_best_cost, d_best_idx, d_workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int mgb_score_host(const mgb_model* m_, const mgb_constraints* c, const void* h_s, int32_t s_is_f64, int64_t n,
                              int64_t index_base, float err_scale, float quality_scale, float magnitude_scale, float* h_cost, int32_t k,
                              float* h_best_cost, int64_t* h_best_idx) {
    MGB_REQUIRE(m_ && c && (h_s || n == 0), "null argument");
    MGB_REQUIRE(k >= 0 && k <= 64, "k must be in [0, 64]");
    MGB_REQUIRE(k == 0 || (h_best_cost && h_best_idx), "k > 0 needs h_best_cost / h_best_idx");
    mgb_model* m = const_cast<mgb_model*>(m_);  // staging buffers are lazily created caches
    std::lock_guard<std::mutex> lock(m->host_mutex);
    MGB_CUDA(cudaSetDevice(m->device));
    const size_t elt = s_is_f64 ? sizeof(double) : sizeof(float);
    const size_t row = (size_t)m->D * elt;
    const int64_t CH = 1 << 18;
    const size_t chunk_bytes = (size_t)CH * row;
    if (!m->copy_stream) {
        MGB_CUDA(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
        MGB_CUDA(cudaStreamCreateWithFlags(&m->compute_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            MGB_CUDA(cudaEventCreateWithFlags(&m->ev_copied[b], cudaEventDisableTiming));
            MGB_CUDA(cudaEventCreateWithFlags(&m->ev_done[b], cudaEventDisableTiming));
        }
    }
    if (m->stage_bytes < chunk_bytes) {
        // both chunks may still be in flight from the previous call's tail only if it failed: drain first
        MGB_CUDA(cudaStreamSynchronize(m->compute_stream));
        for (int b = 0; b < 2; ++b) {
            cudaFree(m->d_stage[b]);
            m->d_stage[b] = nullptr;
        }
        m->stage_bytes = 0;
        for (int b = 0; b < 2; ++b) MGB_CUDA(cudaMalloc(&m->d_stage[b], chunk_bytes));
        m->stage_bytes = chunk_bytes;
    }
    const int64_t n_chunks = n > 0 ? (n + CH - 1) / CH : 0;
    const bool need_all_costs = (h_cost != nullptr) || (k > 1);
    // workspace layout: [score ws][chunk best cost][chunk best idx][costs (optional)]
    const size_t ws_score = mgb_score_workspace_bytes(n, k > 1 ? k : 1);   // k > 1: mgb_topk keeps its stage-1 partials in it
    const size_t best_off = (ws_score + 255) & ~(size_t)255;
    const size_t n_best = (size_t)std::max<int64_t>(n_chunks, 64);
    const size_t idx_off = (best_off + n_best * sizeof(float) + 255) & ~(size_t)255;
    const size_t cost_off = (idx_off + n_best * sizeof(int64_t) + 255) & ~(size_t)255;
    const size_t need = cost_off + (need_all_costs ? (size_t)n * sizeof(float) : 0);
    if (m->ws_bytes < need) {
        cudaFree(m->d_ws);
        m->d_ws = nullptr;
        m->ws_bytes = 0;
        MGB_CUDA(cudaMalloc(&m->d_ws, need));
        MGB_CUDA(cudaMemset(m->d_ws, 0, need));
        m->ws_bytes = need;
    }
    unsigned char* ws = reinterpret_cast<unsigned char*>(m->d_ws);
    float* d_bc = reinterpret_cast<float*>(ws + best_off);
    int64_t* d_bi = reinterpret_cast<int64_t*>(ws + idx_off);
    float* d_cost = need_all_costs ? reinterpret_cast<float*>(ws + cost_off) : nullptr;

    cudaPointerAttributes attr;
    bool host_pinned = false;
    if (n > 0 && cudaPointerGetAttributes(&attr, h_s) == cudaSuccess) host_pinned = (attr.type == cudaMemoryTypeHost);
    cudaGetLastError();  // pageable memory reports an error on old drivers; clear it
    if (!host_pinned && n > 0 && m->pinned_bytes < chunk_bytes) {
        // sized separately from d_stage: a pinned-input call may have grown d_stage without ever allocating these
        for (int b = 0; b < 2; ++b) {
            if (m->pinned[b]) cudaFreeHost(m->pinned[b]);
            m->pinned[b] = nullptr;
        }
        m->pinned_bytes = 0;
        for (int b = 0; b < 2; ++b) MGB_CUDA(cudaMallocHost(&m->pinned[b], chunk_bytes));
        m->pinned_bytes = chunk_bytes;
    }
    for (int64_t ch = 0; ch < n_chunks; ++ch) {
        const int b = (int)(ch & 1);
        const int64_t lo = ch * CH, cnt = std::min<int64_t>(CH, n - lo);
        const unsigned char* src = reinterpret_cast<const unsigned char*>(h_s) + (size_t)lo * row;
        if (ch >= 2) MGB_CUDA(cudaEventSynchronize(m->ev_done[b]));
        if (host_pinned) {
            MGB_CUDA(cudaMemcpyAsync(m->d_stage[b], src, (size_t)cnt * row, cudaMemcpyHostToDevice, m->copy_stream));
        } else {
            memcpy(m->pinned[b], src, (size_t)cnt * row);
            MGB_CUDA(cudaMemcpyAsync(m->d_stage[b], m->pinned[b], (size_t)cnt * row, cudaMemcpyHostToDevice, m->copy_stream));
        }
        MGB_CUDA(cudaEventRecord(m->ev_copied[b], m->copy_stream));
        MGB_CUDA(cudaStreamWaitEvent(m->compute_stream, m->ev_copied[b], 0));
        int rc = mgb_launch_score(m, c, 0, m->d_stage[b], s_is_f64, nullptr, nullptr, 0, cnt, index_base + lo, err_scale, quality_scale,
                                  magnitude_scale, d_cost ? d_cost + lo : nullptr, nullptr, (k >= 1) ? 1 : 0, d_bc + ch, d_bi + ch, ws,
                                  ws_score, m->compute_stream);
        if (rc != MGB_OK) return rc;
        MGB_CUDA(cudaEventRecord(m->ev_done[b], m->compute_stream));
    }
    if (h_cost && n > 0) MGB_CUDA(cudaMemcpyAsync(h_cost, d_cost, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, m->compute_stream));
    if (k == 1) {
        std::vector<float> bc(std::max<int64_t>(n_chunks, 1));
        std::vector<int64_t> bi(std::max<int64_t>(n_chunks, 1));
        if (n_chunks > 0) {
            MGB_CUDA(cudaMemcpyAsync(bc.data(), d_bc, n_chunks * sizeof(float), cudaMemcpyDeviceToHost, m->compute_stream));
            MGB_CUDA(cudaMemcpyAsync(bi.data(), d_bi, n_chunks * sizeof(int64_t), cudaMemcpyDeviceToHost, m->compute_stream));
        }
        MGB_CUDA(cudaStreamSynchronize(m->compute_stream));
        float best = INFINITY;
        int64_t besti = index_base;
        bool have = false;
        for (int64_t ch = 0; ch < n_chunks; ++ch) {
            const float cc = bc[ch];
            if (cc != cc) continue;  // NaN never wins (reference: `min_error > error`)
            if (!have || cc < best || (cc == best && bi[ch] < besti)) {
                best = cc;
                besti = bi[ch];
                have = true;
            }
        }
        h_best_cost[0] = best;
        h_best_idx[0] = besti;
    } else if (k > 1) {
        int rc = mgb_topk(d_cost, n, index_base, k, d_bc, d_bi, ws, ws_score, m->compute_stream);
        if (rc != MGB_OK) return rc;
        MGB_CUDA(cudaMemcpyAsync(h_best_cost, d_bc, k * sizeof(float), cudaMemcpyDeviceToHost, m->compute_stream));
        MGB_CUDA(cudaMemcpyAsync(h_best_idx, d_bi, k * sizeof(int64_t), cudaMemcpyDeviceToHost, m->compute_stream));
        MGB_CUDA(cudaStreamSynchronize(m->compute_stream));
    } else {
        MGB_CUDA(cudaStreamSynchronize(m->compute_stream));
    }
    return MGB_OK;
}

