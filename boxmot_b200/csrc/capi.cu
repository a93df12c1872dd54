// extern "C" surface of libboxmot_b200.so (declarations: include/boxmot_b200.h).
#include "../../include/boxmot_b200.h"
#include "ocsort_engine.cuh"
#include "bytetrack_engine.cuh"
#include "boosttrack_engine.cuh"
#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <vector>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess) snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_err, sizeof(g_err), "%s", what);
    return code;
}

#define BM_CUDA(call)                                         \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return fail(-2, #call, e_);    \
    } while (0)

// Independent groups of streams stepped on their own CUDA streams: fork from the caller's stream,
// one chain of launches per group, join back.  The per-stream phases of a frame are short and
// dependent (predict -> associate -> update -> finalize), and the association kernel ends with its
// slowest CTA; with groups, one group's streaming kernels run under another group's tail.
struct StreamGroups {
    int n = 1;
    std::vector<cudaStream_t> streams;
    std::vector<cudaEvent_t> done;
    cudaEvent_t fork = nullptr;
    int set(int groups, int S) {
        clear();
        groups = groups < 1 ? 1 : (groups > S ? (S > 0 ? S : 1) : groups);
        if (groups > 1) {
            if (cudaEventCreateWithFlags(&fork, cudaEventDisableTiming) != cudaSuccess) return -1;
            for (int g = 0; g < groups; g++) {
                cudaStream_t st; cudaEvent_t ev;
                if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) return -1;
                streams.push_back(st);
                if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return -1;
                done.push_back(ev);
            }
        }
        n = groups;
        return 0;
    }
    void clear() {
        for (cudaStream_t st : streams) cudaStreamDestroy(st);
        for (cudaEvent_t ev : done) cudaEventDestroy(ev);
        if (fork) cudaEventDestroy(fork);
        streams.clear(); done.clear(); fork = nullptr; n = 1;
    }
    // streams [first, first + count) of group g out of S
    static void span(int g, int n, int S, int* first, int* count) {
        const int base = S / n, rem = S % n;
        *first = g * base + (g < rem ? g : rem);
        *count = base + (g < rem ? 1 : 0);
    }
};

struct OcsEngine {
    bm::OcsState st;
    bm::OcsParams pr;
    std::vector<void*> allocs;
    size_t smem;
    cudaEvent_t ev[7];
    bool timing = false, timed = false;
    bool deep = false;
    StreamGroups grp;
};

struct ByteEngine {      // ByteTrack and BoT-SORT share the state layout and the step kernel
    bm::ByteState st;
    bm::ByteParams pr;
    std::vector<void*> allocs;
    bool bot = false;
    StreamGroups grp;
};

struct BoostEngine {
    bm::BoostState st;
    bm::BoostParams pr;
    std::vector<void*> allocs;
    cudaEvent_t ev[7];
    bool timing = false, timed = false;
    StreamGroups grp;
    int* pc_frame = nullptr;   // [S] frame counter at the frame's start (bm_boosttrack_hold_frame)
};

// per_class_decorator (boxmot/trackers/basetracker.py:154-201) for the BYTE family: before class `cls` is updated
// the decorator makes per_class_active_tracks[cls] the tracker's active_tracks and resets frame_count to the
// frame's starting value; lost_stracks / removed_stracks are NOT swapped (one list for all classes).
__global__ void byte_select_class_kernel(bm::ByteState st, int* pc_act, int* pc_n, int* pc_meta, int C, int cls,
                                         int new_frame) {
    const int s = blockIdx.x, tid = threadIdx.x;
    int* hdr = st.hdr + (size_t)s * bm::BH_COUNT;
    int* act = st.act + (size_t)s * st.Tcap;
    int* meta = pc_meta + (size_t)s * 2;
    const int cur = meta[0], n = hdr[bm::BH_NACT];
    if (cur >= 0) {
        int* dst = pc_act + ((size_t)s * C + cur) * st.Tcap;
        for (int i = tid; i < n; i += blockDim.x) dst[i] = act[i];
        if (tid == 0) pc_n[(size_t)s * C + cur] = n;
    }
    __syncthreads();
    const int m = pc_n[(size_t)s * C + cls];
    const int* src = pc_act + ((size_t)s * C + cls) * st.Tcap;
    for (int i = tid; i < m; i += blockDim.x) act[i] = src[i];
    __syncthreads();
    if (tid == 0) {
        hdr[bm::BH_NACT] = m;
        if (new_frame) meta[1] = hdr[bm::BH_FRAME]; else hdr[bm::BH_FRAME] = meta[1];
        meta[0] = cls;
    }
}

__global__ void boost_hold_frame_kernel(bm::BoostState st, int* pc_frame, int new_frame) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= st.S) return;
    int* hdr = st.hdr + (size_t)s * bm::BTH_COUNT;
    if (new_frame) pc_frame[s] = hdr[bm::BTH_FRAME]; else hdr[bm::BTH_FRAME] = pc_frame[s];
}

template <typename E, typename T>
cudaError_t dev_alloc(E* e, T** p, size_t n) {
    cudaError_t r = cudaMalloc(reinterpret_cast<void**>(p), n * sizeof(T));
    if (r == cudaSuccess) e->allocs.push_back(*p);
    return r;
}

}  // namespace

// Lets the other translation units of the library leave a message for bm_last_error().
namespace bm {
void set_last_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }
}  // namespace bm

extern "C" {

const char* bm_last_error(void) { return g_err; }
int bm_version(void) { return 100; }

static int ocs_create(const bm_ocsort_config* cfg, const bm_deepocsort_config* deep, int n_streams, int track_cap,
                      int det_cap, int smem_cost_bytes, void** handle) {
    if (!cfg || !handle || n_streams <= 0 || track_cap <= 0 || det_cap <= 0)
        return fail(-1, "bm_ocsort_create: bad argument");
    if (cfg->delta_t < 1 || cfg->delta_t > 16) return fail(-1, "bm_ocsort_create: delta_t out of range");
    if (track_cap > 1024) return fail(-1, "bm_ocsort_create: track_cap must be <= 1024");
    OcsEngine* e = new (std::nothrow) OcsEngine();
    if (!e) return fail(-3, "out of host memory");
    bm::OcsState& st = e->st;
    st.prof = nullptr;
    st.S = n_streams; st.Tcap = track_cap; st.Dcap = det_cap; st.R = cfg->delta_t;
    // DeepOCSORT keeps the masked embedding matrix next to the cost matrix
    st.scratch_elems = (size_t)track_cap * det_cap * (deep ? 2 : 1);
    st.smem_cost_elems = smem_cost_bytes / 8;
    st.emb = nullptr; st.embsim = nullptr; st.bornslot = nullptr; st.epos = nullptr; st.wbuf = nullptr; st.dP = nullptr;
    st.deep_d = nullptr; st.deep_i = nullptr;
    e->deep = deep != nullptr;
    bm::OcsParams& pr = e->pr;
    pr.det_thresh = cfg->det_thresh; pr.min_conf = cfg->min_conf; pr.iou_thr = cfg->asso_threshold;
    pr.inertia = cfg->inertia; pr.q_xy = cfg->q_xy_scaling; pr.q_s = cfg->q_s_scaling;
    pr.diag = sqrt((double)cfg->img_w * cfg->img_w + (double)cfg->img_h * cfg->img_h);
    pr.max_age = cfg->max_age; pr.min_hits = cfg->min_hits; pr.delta_t = cfg->delta_t;
    pr.asso_func = cfg->asso_func; pr.use_byte = cfg->use_byte ? 1 : 0;
    pr.emb_on = 0; pr.aw_off = 1; pr.E = 0; pr.w_emb = pr.alpha_fixed = pr.aw_param = 0.0;
    // The separate screen launch shortens every association CTA (mean lifetime 157 k -> 89 k cycles) but the
    // kernel's span at up to one wave of CTAs (8 x 148 streams) is its slowest solver chain, so it only
    // pays once the batch is throughput-bound: measured +3..8 % at 2048 / 4096 streams, -0..4 % at <= 1184.
    // BM_PRESCREEN=0/1 overrides.
    pr.prescreen = n_streams > 1536 ? 1 : 0;
    if (const char* ps = getenv("BM_PRESCREEN")) pr.prescreen = ps[0] == '1' ? 1 : 0;
    const size_t S = n_streams, T = track_cap;
    cudaError_t r = cudaSuccess;
    if (deep) {
        pr.emb_on = deep->embedding_off ? 0 : 1; pr.aw_off = deep->aw_off ? 1 : 0; pr.E = deep->emb_dim;
        pr.w_emb = deep->w_association_emb; pr.alpha_fixed = deep->alpha_fixed_emb; pr.aw_param = deep->aw_param;
        if (pr.emb_on) {
            if (r == cudaSuccess) r = dev_alloc(e, &st.emb, S * T * (size_t)pr.E);
            if (r == cudaSuccess) r = dev_alloc(e, &st.embsim, S * (size_t)det_cap * T);
            if (r == cudaSuccess) r = dev_alloc(e, &st.bornslot, S * (size_t)det_cap);
        }
        if (r == cudaSuccess) r = dev_alloc(e, &st.epos, S * T);
        if (r == cudaSuccess) r = dev_alloc(e, &st.wbuf, S * ((size_t)det_cap + T));
        if (r == cudaSuccess) r = dev_alloc(e, &st.dP, S * 98 * T);      // dense covariances (tracks that saw a rotating warp)
        if (r == cudaSuccess) r = dev_alloc(e, &st.deep_d, S * ((size_t)det_cap + 1));
        if (r == cudaSuccess) r = dev_alloc(e, &st.deep_i, S * (3 * (size_t)det_cap + T));
    }
    if (r == cudaSuccess) r = dev_alloc(e, &st.xP, S * bm::XB_N * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.sxP, S * bm::XB_N * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.lastz, S * 4 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.lastobs, S * 5 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.vel, S * 2 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.confcls, S * 2 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.ring, S * st.R * 4 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.ring_age, S * st.R * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.ints, S * bm::OI_COUNT * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.order, S * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.hdr, S * bm::OH_COUNT);
    if (r == cudaSuccess) r = dev_alloc(e, &st.n_updates, S);
    if (r == cudaSuccess) r = dev_alloc(e, &st.scratch, S * st.scratch_elems);
    if (r == cudaSuccess) r = dev_alloc(e, &st.abuf, S * bm::AB_COUNT * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.upd, S * 2 * (size_t)det_cap);
    if (r == cudaSuccess) r = dev_alloc(e, &st.born, S * (size_t)det_cap);
    if (r == cudaSuccess) r = dev_alloc(e, &st.hsel, S * 2 * (size_t)det_cap);
    if (r == cudaSuccess) r = dev_alloc(e, &st.scan, S * (size_t)det_cap);
    e->smem = bm::ocsort_step_smem(st, pr);
    if (r == cudaSuccess) r = bm::ocsort_step_configure(e->smem);
    if (r != cudaSuccess) {
        for (void* p : e->allocs) cudaFree(p);
        delete e;
        return fail(-2, "bm_ocsort_create", r);
    }
    *handle = e;
    return bm_ocsort_reset(e, nullptr);
}

int bm_ocsort_create(const bm_ocsort_config* cfg, int n_streams, int track_cap, int det_cap,
                     int smem_cost_bytes, void** handle) {
    return ocs_create(cfg, nullptr, n_streams, track_cap, det_cap, smem_cost_bytes, handle);
}

int bm_deepocsort_create(const bm_deepocsort_config* cfg, int n_streams, int track_cap, int det_cap, void** handle) {
    if (!cfg) return fail(-1, "bm_deepocsort_create: bad argument");
    if (!cfg->embedding_off && cfg->emb_dim <= 0) return fail(-1, "bm_deepocsort_create: emb_dim must be > 0");
    if (cfg->oc.use_byte) return fail(-1, "bm_deepocsort_create: DeepOCSORT has no BYTE round");
    if (!(cfg->aw_param < 1.0)) return fail(-1, "bm_deepocsort_create: aw_param must be < 1");
    bm_ocsort_config oc = cfg->oc;
    oc.min_conf = oc.det_thresh;          // one confidence gate: scores > det_thresh (deepocsort.py:340-343)
    return ocs_create(&oc, cfg, n_streams, track_cap, det_cap, 0, handle);
}

int bm_deepocsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !d_dets || !d_ndets || !d_out || !d_nout) return fail(-1, "bm_deepocsort_step: null argument");
    if (!e->deep) return fail(-1, "bm_deepocsort_step: not a DeepOCSORT engine");
    if (e->pr.emb_on && !d_embs) return fail(-1, "bm_deepocsort_step: embeddings required unless embedding_off");
    cudaStream_t main = static_cast<cudaStream_t>(stream);
    if (e->grp.n > 1 && !e->timing) {
        const bm::OcsState& st = e->st;
        const size_t D = st.Dcap, orow = (size_t)(st.Tcap < st.Dcap ? st.Tcap : st.Dcap) * 8;
        BM_CUDA(cudaEventRecord(e->grp.fork, main));
        for (int g = 0; g < e->grp.n; g++) {
            int s0, n;
            StreamGroups::span(g, e->grp.n, st.S, &s0, &n);
            cudaStream_t gs = e->grp.streams[g];
            BM_CUDA(cudaStreamWaitEvent(gs, e->grp.fork, 0));
            BM_CUDA(bm::deepocsort_step_launch(bm::ocs_substate(st, s0, n, e->pr.E), e->pr, d_dets + s0 * D * 6,
                                               d_ndets + s0, d_embs ? d_embs + s0 * D * e->pr.E : nullptr,
                                               d_warps ? d_warps + (size_t)s0 * 6 : nullptr, d_out + s0 * orow,
                                               d_nout + s0, gs, nullptr));
            BM_CUDA(cudaEventRecord(e->grp.done[g], gs));
            BM_CUDA(cudaStreamWaitEvent(main, e->grp.done[g], 0));
        }
        e->timed = false;
        return 0;
    }
    BM_CUDA(bm::deepocsort_step_launch(e->st, e->pr, d_dets, d_ndets, d_embs, d_warps, d_out, d_nout, main,
                                       e->timing ? e->ev : nullptr));
    e->timed = e->timing;
    return 0;
}

int bm_deepocsort_kernel_times(void* handle, float* ms6) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !ms6 || !e->deep) return fail(-1, "bm_deepocsort_kernel_times: bad argument");
    if (!e->timed) return fail(-1, "bm_deepocsort_kernel_times: no timed step (call bm_ocsort_time_kernels first)");
    BM_CUDA(cudaEventSynchronize(e->ev[6]));
    for (int i = 0; i < 6; i++) BM_CUDA(cudaEventElapsedTime(&ms6[i], e->ev[i], e->ev[i + 1]));
    return 0;
}

int bm_ocsort_destroy(void* handle) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return 0;
    if (e->timing) for (int i = 0; i < 7; i++) cudaEventDestroy(e->ev[i]);
    e->grp.clear();
    for (void* p : e->allocs) cudaFree(p);
    delete e;
    return 0;
}

int bm_ocsort_reset(void* handle, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::OcsState& st = e->st;
    const size_t S = st.S, T = st.Tcap;
    BM_CUDA(cudaMemsetAsync(st.hdr, 0, S * bm::OH_COUNT * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.n_updates, 0, S * sizeof(unsigned long long), s));
    BM_CUDA(cudaMemsetAsync(st.order, 0, S * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.ints, 0, S * bm::OI_COUNT * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.xP, 0, S * bm::XB_N * T * sizeof(double), s));
    BM_CUDA(cudaMemsetAsync(st.sxP, 0, S * bm::XB_N * T * sizeof(double), s));
    if (st.emb) BM_CUDA(cudaMemsetAsync(st.emb, 0, S * T * (size_t)e->pr.E * sizeof(float), s));
    return 0;
}

int bm_ocsort_step(void* handle, const float* d_dets, const int* d_ndets, double* d_out, int* d_nout,
                   void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !d_dets || !d_ndets || !d_out || !d_nout) return fail(-1, "bm_ocsort_step: null argument");
    if (e->deep) return fail(-1, "bm_ocsort_step: DeepOCSORT engine (use bm_deepocsort_step)");
    cudaStream_t main = static_cast<cudaStream_t>(stream);
    if (e->grp.n > 1 && !e->timing) {
        const bm::OcsState& st = e->st;
        const size_t D = st.Dcap, orow = (size_t)(st.Tcap < st.Dcap ? st.Tcap : st.Dcap) * 8;
        BM_CUDA(cudaEventRecord(e->grp.fork, main));
        for (int g = 0; g < e->grp.n; g++) {
            int s0, n;
            StreamGroups::span(g, e->grp.n, st.S, &s0, &n);
            cudaStream_t gs = e->grp.streams[g];
            BM_CUDA(cudaStreamWaitEvent(gs, e->grp.fork, 0));
            BM_CUDA(bm::ocsort_step_launch(bm::ocs_substate(st, s0, n, 0), e->pr, d_dets + s0 * D * 6, d_ndets + s0,
                                           d_out + s0 * orow, d_nout + s0, gs, nullptr));
            BM_CUDA(cudaEventRecord(e->grp.done[g], gs));
            BM_CUDA(cudaStreamWaitEvent(main, e->grp.done[g], 0));
        }
        e->timed = false;
        return 0;
    }
    BM_CUDA(bm::ocsort_step_launch(e->st, e->pr, d_dets, d_ndets, d_out, d_nout, main, e->timing ? e->ev : nullptr));
    e->timed = e->timing;
    return 0;
}

int bm_ocsort_step_frames(void* handle, const float* d_dets, const int* d_ndets, double* d_out, int* d_nout,
                          int n_frames, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !d_dets || !d_ndets || !d_out || !d_nout || n_frames < 0) return fail(-1, "bm_ocsort_step_frames: bad argument");
    if (e->deep) return fail(-1, "bm_ocsort_step_frames: DeepOCSORT engine");
    const bm::OcsState& st = e->st;
    const size_t S = st.S, D = st.Dcap, orow = (size_t)(st.Tcap < st.Dcap ? st.Tcap : st.Dcap) * 8;
    cudaStream_t main = static_cast<cudaStream_t>(stream);
    if (e->grp.n <= 1) {
        for (int f = 0; f < n_frames; f++)
            BM_CUDA(bm::ocsort_step_launch(st, e->pr, d_dets + f * S * D * 6, d_ndets + f * S, d_out + f * S * orow,
                                           d_nout + f * S, main, nullptr));
        e->timed = false;
        return 0;
    }
    // every group runs its chain of launches for ALL frames back to back on its own CUDA stream: a group that
    // is waiting for the slowest of its streams no longer holds the others back frame by frame
    BM_CUDA(cudaEventRecord(e->grp.fork, main));
    for (int g = 0; g < e->grp.n; g++) BM_CUDA(cudaStreamWaitEvent(e->grp.streams[g], e->grp.fork, 0));
    for (int f = 0; f < n_frames; f++)
        for (int g = 0; g < e->grp.n; g++) {
            int s0, n;
            StreamGroups::span(g, e->grp.n, st.S, &s0, &n);
            BM_CUDA(bm::ocsort_step_launch(bm::ocs_substate(st, s0, n, 0), e->pr, d_dets + (f * S + s0) * D * 6,
                                           d_ndets + f * S + s0, d_out + (f * S + s0) * orow, d_nout + f * S + s0,
                                           e->grp.streams[g], nullptr));
        }
    for (int g = 0; g < e->grp.n; g++) {
        BM_CUDA(cudaEventRecord(e->grp.done[g], e->grp.streams[g]));
        BM_CUDA(cudaStreamWaitEvent(main, e->grp.done[g], 0));
    }
    e->timed = false;
    return 0;
}

int bm_ocsort_set_groups(void* handle, int groups) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    if (e->grp.set(groups, e->st.S) != 0) return fail(-2, "bm_ocsort_set_groups: stream / event creation failed");
    return e->grp.n;
}

int bm_ocsort_time_kernels(void* handle, int enable) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    if (enable && !e->timing) {
        for (int i = 0; i < 7; i++) BM_CUDA(cudaEventCreate(&e->ev[i]));
        e->timing = true;
    } else if (!enable && e->timing) {
        for (int i = 0; i < 7; i++) cudaEventDestroy(e->ev[i]);
        e->timing = e->timed = false;
    }
    return 0;
}

int bm_ocsort_kernel_times(void* handle, float* ms4) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !ms4 || e->deep) return fail(-1, "bm_ocsort_kernel_times: bad argument");
    if (!e->timed) return fail(-1, "bm_ocsort_kernel_times: no timed step (call bm_ocsort_time_kernels first)");
    BM_CUDA(cudaEventSynchronize(e->ev[4]));
    for (int i = 0; i < 4; i++) BM_CUDA(cudaEventElapsedTime(&ms4[i], e->ev[i], e->ev[i + 1]));
    return 0;
}

int bm_ocsort_status(void* handle, int* err_bits, unsigned long long* track_updates,
                     unsigned long long* lap_solves, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int S = e->st.S;
    std::vector<int> hdr((size_t)S * bm::OH_COUNT);
    std::vector<unsigned long long> nu(S);
    BM_CUDA(cudaMemcpyAsync(hdr.data(), e->st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(nu.data(), e->st.n_updates, nu.size() * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    int err = 0;
    unsigned long long tot = 0, laps = 0;
    for (int i = 0; i < S; i++) {
        err |= hdr[(size_t)i * bm::OH_COUNT + bm::OH_ERR];
        laps += (unsigned long long)hdr[(size_t)i * bm::OH_COUNT + bm::OH_LAPS];
        tot += nu[i];
    }
    if (err_bits) *err_bits = err;
    if (track_updates) *track_updates = tot;
    if (lap_solves) *lap_solves = laps;
    return 0;
}

int bm_ocsort_debug_profile(void* handle, long long* d_stamps) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    e->st.prof = d_stamps;
    return 0;
}

int bm_ocsort_counters(void* handle, unsigned long long* out4, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || !out4) return fail(-1, "bm_ocsort_counters: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int S = e->st.S;
    std::vector<int> hdr((size_t)S * bm::OH_COUNT);
    std::vector<unsigned long long> nu(S);
    BM_CUDA(cudaMemcpyAsync(hdr.data(), e->st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(nu.data(), e->st.n_updates, nu.size() * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
    for (int i = 0; i < S; i++) {
        out4[0] += nu[i];
        out4[1] += (unsigned long long)hdr[(size_t)i * bm::OH_COUNT + bm::OH_LAPS];
        out4[2] += (unsigned long long)hdr[(size_t)i * bm::OH_COUNT + bm::OH_MATCH];
        out4[3] += (unsigned long long)hdr[(size_t)i * bm::OH_COUNT + bm::OH_BIRTHS];
    }
    return 0;
}

int bm_ocsort_stream_info(void* handle, int* h_n_tracks, int* h_next_id, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int S = e->st.S;
    std::vector<int> hdr((size_t)S * bm::OH_COUNT);
    BM_CUDA(cudaMemcpyAsync(hdr.data(), e->st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    for (int i = 0; i < S; i++) {
        if (h_n_tracks) h_n_tracks[i] = hdr[(size_t)i * bm::OH_COUNT + bm::OH_NLIST];
        if (h_next_id) h_next_id[i] = hdr[(size_t)i * bm::OH_COUNT + bm::OH_NEXT_ID];
    }
    return 0;
}

int bm_ocsort_peek(void* handle, int si, int* n_tracks, double* h_x, double* h_P, int* h_ids, void* stream) {
    OcsEngine* e = static_cast<OcsEngine*>(handle);
    if (!e || si < 0 || si >= e->st.S) return fail(-1, "bm_ocsort_peek: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::OcsState& st = e->st;
    const size_t T = st.Tcap;
    std::vector<double> xp(bm::XB_N * T), dp(st.dP ? 49 * T : 0);
    std::vector<int> order(T), ints(bm::OI_COUNT * T), hdr(bm::OH_COUNT);
    BM_CUDA(cudaMemcpyAsync(xp.data(), st.xP + (size_t)si * bm::XB_N * T, xp.size() * 8, cudaMemcpyDeviceToHost, s));
    if (st.dP) BM_CUDA(cudaMemcpyAsync(dp.data(), st.dP + (size_t)si * 98 * T, dp.size() * 8, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(order.data(), st.order + (size_t)si * T, T * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(ints.data(), st.ints + (size_t)si * bm::OI_COUNT * T, ints.size() * 4,
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(hdr.data(), st.hdr + (size_t)si * bm::OH_COUNT, hdr.size() * 4,
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    const int n = hdr[bm::OH_NLIST];
    if (n_tracks) *n_tracks = n;
    for (int p = 0; p < n; p++) {
        const int slot = order[p];
        if (h_x) for (int c = 0; c < 7; c++) h_x[p * 7 + c] = xp[c * T + slot];
        if (h_P && st.dP && (ints[bm::OI_FLAGS * T + slot] & bm::OF_DENSE)) {
            for (int c = 0; c < 49; c++) h_P[(size_t)p * 49 + c] = dp[c * T + slot];
        } else if (h_P) {
            double pb[13];
            for (int c = 0; c < 13; c++) pb[c] = xp[(7 + c) * T + slot];
            bm::xysrb_expand(pb, h_P + (size_t)p * 49);
        }
        if (h_ids) h_ids[p] = ints[bm::OI_ID * T + slot];
    }
    return 0;
}

// ---------------------------------------------------------------------------- BoostTrack
int bm_boosttrack_create(const bm_boosttrack_config* cfg, int n_streams, int track_cap, int det_cap, void** handle) {
    if (!cfg || !handle || n_streams <= 0 || track_cap <= 0 || det_cap <= 0)
        return fail(-1, "bm_boosttrack_create: bad argument");
    if (track_cap > 1024) return fail(-1, "bm_boosttrack_create: track_cap must be <= 1024");
    if (cfg->with_reid && cfg->emb_dim <= 0) return fail(-1, "bm_boosttrack_create: with_reid needs emb_dim > 0");
    if (cfg->max_age < 0) return fail(-1, "bm_boosttrack_create: max_age must be >= 0");
    BoostEngine* e = new (std::nothrow) BoostEngine();
    if (!e) return fail(-3, "out of host memory");
    bm::BoostState& st = e->st;
    bm::BoostParams& pr = e->pr;
    st = bm::BoostState{};
    st.S = n_streams; st.Tcap = track_cap; st.Dcap = det_cap;
    pr.det_thresh = cfg->det_thresh; pr.iou_thr = cfg->iou_threshold; pr.min_box_area = cfg->min_box_area;
    pr.ar_thresh = cfg->aspect_ratio_thresh; pr.l_iou = cfg->lambda_iou; pr.l_mhd = cfg->lambda_mhd;
    pr.l_shape = cfg->lambda_shape; pr.dlo_coef = cfg->dlo_boost_coef;
    pr.max_age = cfg->max_age; pr.min_hits = cfg->min_hits; pr.use_dlo = cfg->use_dlo_boost ? 1 : 0;
    pr.use_duo = cfg->use_duo_boost ? 1 : 0; pr.s_corr = cfg->s_sim_corr ? 1 : 0; pr.rich_s = cfg->use_rich_s ? 1 : 0;
    pr.use_sb = cfg->use_sb ? 1 : 0; pr.use_vt = cfg->use_vt ? 1 : 0; pr.with_reid = cfg->with_reid ? 1 : 0;
    pr.E = cfg->with_reid ? cfg->emb_dim : 0;
    const size_t S = n_streams, T = track_cap, D = det_cap;
    // KalmanBoxTracker.get_confidence: 0.9 ** k through the host's libm, as Python evaluates it
    st.npow = (cfg->max_age + 2 > 8 ? cfg->max_age + 2 : 8);
    std::vector<double> pw(st.npow);
    for (int k = 0; k < st.npow; k++) pw[k] = std::pow(0.9, (double)k);
    double* d_pw = nullptr;
    cudaError_t r = cudaSuccess;
    if (r == cudaSuccess) r = dev_alloc(e, &d_pw, (size_t)st.npow);
    if (r == cudaSuccess) r = cudaMemcpy(d_pw, pw.data(), pw.size() * sizeof(double), cudaMemcpyHostToDevice);
    st.powtab = d_pw;
    if (r == cudaSuccess) r = dev_alloc(e, &st.xP, S * bm::BTX_N * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.confcls, S * 2 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.ints, S * bm::BTI_COUNT * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.order, S * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.hdr, S * bm::BTH_COUNT);
    if (r == cudaSuccess) r = dev_alloc(e, &st.n_updates, S);
    if (r == cudaSuccess) r = dev_alloc(e, &st.tp, S * bm::TP_COUNT * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.dconf, S * D);
    if (r == cudaSuccess) r = dev_alloc(e, &st.hsel, S * D);
    st.cost_stride = std::max(D * T, bm::bt_list_doubles(track_cap));
    if (r == cudaSuccess) r = dev_alloc(e, &st.cost, S * st.cost_stride);
    if (r == cudaSuccess) r = dev_alloc(e, &st.upd, S * 2 * D);
    if (r == cudaSuccess) r = dev_alloc(e, &st.born, S * D);
    if (r == cudaSuccess) r = dev_alloc(e, &st.embsim, S * D * T);      // read (as zeros) only with_reid
    if (pr.with_reid) {
        if (r == cudaSuccess) r = dev_alloc(e, &st.bornslot, S * D);
        if (r == cudaSuccess) r = dev_alloc(e, &st.emb, S * T * (size_t)pr.E);
    }
    if (r == cudaSuccess) r = bm::boosttrack_configure(st);
    if (r != cudaSuccess) {
        for (void* p : e->allocs) cudaFree(p);
        delete e;
        return fail(-2, "bm_boosttrack_create", r);
    }
    *handle = e;
    return bm_boosttrack_reset(e, nullptr);
}

int bm_boosttrack_destroy(void* handle) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e) return 0;
    if (e->timing) for (int i = 0; i < 7; i++) cudaEventDestroy(e->ev[i]);
    e->grp.clear();
    for (void* p : e->allocs) cudaFree(p);
    delete e;
    return 0;
}

int bm_boosttrack_reset(void* handle, void* stream) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::BoostState& st = e->st;
    const size_t S = st.S, T = st.Tcap;
    BM_CUDA(cudaMemsetAsync(st.hdr, 0, S * bm::BTH_COUNT * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.n_updates, 0, S * sizeof(unsigned long long), s));
    BM_CUDA(cudaMemsetAsync(st.order, 0, S * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.ints, 0, S * bm::BTI_COUNT * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.xP, 0, S * bm::BTX_N * T * sizeof(double), s));
    if (st.emb) BM_CUDA(cudaMemsetAsync(st.emb, 0, S * T * (size_t)e->pr.E * sizeof(float), s));
    return 0;
}

int bm_boosttrack_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e || !d_dets || !d_ndets || !d_out || !d_nout) return fail(-1, "bm_boosttrack_step: null argument");
    if (e->pr.with_reid && !d_embs) return fail(-1, "bm_boosttrack_step: with_reid needs embeddings");
    cudaStream_t main = static_cast<cudaStream_t>(stream);
    if (e->grp.n > 1 && !e->timing) {
        const bm::BoostState& st = e->st;
        const size_t D = st.Dcap, orow = (size_t)(st.Tcap < st.Dcap ? st.Tcap : st.Dcap) * 8;
        BM_CUDA(cudaEventRecord(e->grp.fork, main));
        for (int g = 0; g < e->grp.n; g++) {
            int s0, n;
            StreamGroups::span(g, e->grp.n, st.S, &s0, &n);
            cudaStream_t gs = e->grp.streams[g];
            BM_CUDA(cudaStreamWaitEvent(gs, e->grp.fork, 0));
            BM_CUDA(bm::boosttrack_step_launch(bm::bt_substate(st, s0, n, e->pr.E), e->pr, d_dets + s0 * D * 6,
                                               d_ndets + s0, d_embs ? d_embs + s0 * D * e->pr.E : nullptr,
                                               d_warps ? d_warps + (size_t)s0 * 6 : nullptr, d_out + s0 * orow,
                                               d_nout + s0, gs, nullptr));
            BM_CUDA(cudaEventRecord(e->grp.done[g], gs));
            BM_CUDA(cudaStreamWaitEvent(main, e->grp.done[g], 0));
        }
        e->timed = false;
        return 0;
    }
    BM_CUDA(bm::boosttrack_step_launch(e->st, e->pr, d_dets, d_ndets, d_embs, d_warps, d_out, d_nout, main,
                                       e->timing ? e->ev : nullptr));
    e->timed = e->timing;
    return 0;
}

int bm_boosttrack_hold_frame(void* handle, int new_frame, void* stream) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e) return fail(-1, "bm_boosttrack_hold_frame: null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!e->pc_frame) {
        BM_CUDA(dev_alloc(e, &e->pc_frame, (size_t)e->st.S));
        BM_CUDA(cudaMemsetAsync(e->pc_frame, 0, (size_t)e->st.S * sizeof(int), s));
    }
    boost_hold_frame_kernel<<<(e->st.S + 127) / 128, 128, 0, s>>>(e->st, e->pc_frame, new_frame);
    BM_CUDA(cudaGetLastError());
    return 0;
}

int bm_boosttrack_set_groups(void* handle, int groups) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    if (e->grp.set(groups, e->st.S) != 0) return fail(-2, "bm_boosttrack_set_groups: stream / event creation failed");
    return e->grp.n;
}

int bm_boosttrack_time_kernels(void* handle, int enable) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    if (enable && !e->timing) {
        for (int i = 0; i < 7; i++) BM_CUDA(cudaEventCreate(&e->ev[i]));
        e->timing = true;
    } else if (!enable && e->timing) {
        for (int i = 0; i < 7; i++) cudaEventDestroy(e->ev[i]);
        e->timing = e->timed = false;
    }
    return 0;
}

int bm_boosttrack_kernel_times(void* handle, float* ms6) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e || !ms6) return fail(-1, "bm_boosttrack_kernel_times: bad argument");
    if (!e->timed) return fail(-1, "bm_boosttrack_kernel_times: no timed step");
    BM_CUDA(cudaEventSynchronize(e->ev[6]));
    for (int i = 0; i < 6; i++) BM_CUDA(cudaEventElapsedTime(&ms6[i], e->ev[i], e->ev[i + 1]));
    return 0;
}

int bm_boosttrack_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e || !out4) return fail(-1, "bm_boosttrack_counters: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int S = e->st.S;
    std::vector<int> hdr((size_t)S * bm::BTH_COUNT);
    std::vector<unsigned long long> nu(S);
    BM_CUDA(cudaMemcpyAsync(hdr.data(), e->st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(nu.data(), e->st.n_updates, nu.size() * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    int err = 0;
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
    for (int i = 0; i < S; i++) {
        err |= hdr[(size_t)i * bm::BTH_COUNT + bm::BTH_ERR];
        out4[0] += nu[i];
        out4[1] += (unsigned long long)hdr[(size_t)i * bm::BTH_COUNT + bm::BTH_LAPS];
        out4[2] += (unsigned long long)hdr[(size_t)i * bm::BTH_COUNT + bm::BTH_MATCH];
        out4[3] += (unsigned long long)hdr[(size_t)i * bm::BTH_COUNT + bm::BTH_BIRTHS];
    }
    if (err_bits) *err_bits = err;
    return 0;
}

int bm_boosttrack_peek(void* handle, int si, int* n_tracks, double* h_x, double* h_P, int* h_ids, void* stream) {
    BoostEngine* e = static_cast<BoostEngine*>(handle);
    if (!e || si < 0 || si >= e->st.S) return fail(-1, "bm_boosttrack_peek: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::BoostState& st = e->st;
    const size_t T = st.Tcap;
    std::vector<double> xp(bm::BTX_N * T);
    std::vector<int> order(T), ints(bm::BTI_COUNT * T), hdr(bm::BTH_COUNT);
    BM_CUDA(cudaMemcpyAsync(xp.data(), st.xP + (size_t)si * bm::BTX_N * T, xp.size() * 8, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(order.data(), st.order + (size_t)si * T, T * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(ints.data(), st.ints + (size_t)si * bm::BTI_COUNT * T, ints.size() * 4,
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(hdr.data(), st.hdr + (size_t)si * bm::BTH_COUNT, hdr.size() * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    const int n = hdr[bm::BTH_NLIST];
    if (n_tracks) *n_tracks = n;
    for (int p = 0; p < n; p++) {
        const int slot = order[p];
        if (h_x) for (int c = 0; c < 8; c++) h_x[p * 8 + c] = xp[c * T + slot];
        if (h_P) {
            double pb[16];
            for (int c = 0; c < 16; c++) pb[c] = xp[(8 + c) * T + slot];
            bm::bt_expand(pb, h_P + (size_t)p * 64);
        }
        if (h_ids) h_ids[p] = ints[bm::BTI_ID * T + slot];
    }
    return 0;
}

// ---------------------------------------------------------------------------- ByteTrack

static int byte_family_create(ByteEngine* e, int n_streams, int track_cap, int det_cap, void** handle) {
    bm::ByteState& st = e->st;
    st.S = n_streams; st.Tcap = track_cap; st.Dcap = det_cap;
    st.feat = st.dfeat = st.embd = nullptr; st.chist = nullptr; st.fupd = nullptr; st.dP = nullptr;
    st.elive = st.ehind = nullptr; st.etn = st.edn = nullptr;
    {   // BM_COSINE=ffma selects the fp32 FFMA contraction (debug / comparison); default: tensor cores
        const char* v = getenv("BM_COSINE");
        st.tensor_cores = (v && v[0] == 'f') ? 0 : 1;
    }
    const size_t S = n_streams, T = track_cap, E = e->pr.emb_dim;
    cudaError_t r = cudaSuccess;
    if (e->bot) {
        if (r == cudaSuccess) r = dev_alloc(e, &st.chist, S * 2 * bm::BOT_HIST * T);
        if (r == cudaSuccess) r = dev_alloc(e, &st.dP, S * 64 * T);       // dense covariances (tracks that saw a rotating warp)
        if (r == cudaSuccess) r = dev_alloc(e, &st.fupd, S * 2 * (size_t)det_cap);
        if (e->pr.with_reid) {
            if (r == cudaSuccess) r = dev_alloc(e, &st.feat, S * T * E);
            if (r == cudaSuccess) r = dev_alloc(e, &st.dfeat, S * (size_t)det_cap * E);
            if (r == cudaSuccess) r = dev_alloc(e, &st.embd, S * T * (size_t)det_cap);
            if (r == cudaSuccess) r = dev_alloc(e, &st.elive, S * T);
            if (r == cudaSuccess) r = dev_alloc(e, &st.ehind, S * (size_t)det_cap);
            if (r == cudaSuccess) r = dev_alloc(e, &st.etn, S * T);
            if (r == cudaSuccess) r = dev_alloc(e, &st.edn, S * (size_t)det_cap);
            if (r == cudaSuccess) r = cudaMemset(st.feat, 0, S * T * E * sizeof(float));
        }
    }
    if (r == cudaSuccess) r = dev_alloc(e, &st.xP, S * bm::XA_N * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.confcls, S * 2 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.tbox, S * 4 * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.ints, S * bm::BI_COUNT * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.act, S * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.lost, S * T);
    if (r == cudaSuccess) r = dev_alloc(e, &st.hdr, S * bm::BH_COUNT);
    if (r == cudaSuccess) r = dev_alloc(e, &st.n_updates, S);
    if (r == cudaSuccess) r = dev_alloc(e, &st.scratch, S * T * (size_t)det_cap);
    if (r == cudaSuccess) r = bm::bytetrack_step_configure(bm::bytetrack_step_smem(st));
    if (r != cudaSuccess) {
        for (void* p : e->allocs) cudaFree(p);
        delete e;
        return fail(-2, "engine create", r);
    }
    *handle = e;
    return bm_bytetrack_reset(e, nullptr);
}

int bm_bytetrack_create(const bm_bytetrack_config* cfg, int n_streams, int track_cap, int det_cap, void** handle) {
    if (!cfg || !handle || n_streams <= 0 || track_cap <= 0 || det_cap <= 0)
        return fail(-1, "bm_bytetrack_create: bad argument");
    if (track_cap > 1024) return fail(-1, "bm_bytetrack_create: track_cap must be <= 1024");
    ByteEngine* e = new (std::nothrow) ByteEngine();
    if (!e) return fail(-3, "out of host memory");
    bm::ByteParams& pr = e->pr;
    pr = bm::ByteParams{};
    pr.track_thresh = cfg->track_thresh; pr.min_conf = cfg->min_conf; pr.match_thresh = cfg->match_thresh;
    pr.det_thresh = cfg->track_thresh;                                   // bytetrack.py:152
    pr.max_time_lost = (int)(cfg->frame_rate / 30.0 * cfg->track_buffer);   // bytetrack.py:153-154
    return byte_family_create(e, n_streams, track_cap, det_cap, handle);
}

int bm_botsort_create(const bm_botsort_config* cfg, int n_streams, int track_cap, int det_cap, void** handle) {
    if (!cfg || !handle || n_streams <= 0 || track_cap <= 0 || det_cap <= 0)
        return fail(-1, "bm_botsort_create: bad argument");
    if (track_cap > 1024) return fail(-1, "bm_botsort_create: track_cap must be <= 1024");
    if (cfg->with_reid && cfg->emb_dim <= 0) return fail(-1, "bm_botsort_create: with_reid needs emb_dim > 0");
    ByteEngine* e = new (std::nothrow) ByteEngine();
    if (!e) return fail(-3, "out of host memory");
    e->bot = true;
    bm::ByteParams& pr = e->pr;
    pr = bm::ByteParams{};
    pr.track_thresh = cfg->track_high_thresh; pr.min_conf = cfg->track_low_thresh;
    pr.match_thresh = cfg->match_thresh; pr.det_thresh = cfg->new_track_thresh;
    pr.max_time_lost = (int)(cfg->frame_rate / 30.0 * cfg->track_buffer);   // botsort.py:76-77
    pr.proximity_thresh = cfg->proximity_thresh; pr.appearance_thresh = cfg->appearance_thresh;
    pr.fuse_first = cfg->fuse_first_associate ? 1 : 0; pr.with_reid = cfg->with_reid ? 1 : 0;
    pr.emb_dim = cfg->with_reid ? cfg->emb_dim : 0;
    return byte_family_create(e, n_streams, track_cap, det_cap, handle);
}

int bm_botsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                    const double* d_warps, double* d_out, int* d_nout, void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || !e->bot || !d_dets || !d_ndets || !d_out || !d_nout) return fail(-1, "bm_botsort_step: bad argument");
    if (e->pr.with_reid && !d_embs) return fail(-1, "bm_botsort_step: with_reid needs embeddings");
    cudaStream_t main = static_cast<cudaStream_t>(stream);
    if (e->grp.n > 1) {
        const bm::ByteState& st = e->st;
        const size_t D = st.Dcap, orow = (size_t)(st.Tcap < st.Dcap ? st.Tcap : st.Dcap) * 8, E = e->pr.emb_dim;
        BM_CUDA(cudaEventRecord(e->grp.fork, main));
        for (int g = 0; g < e->grp.n; g++) {
            int s0, n;
            StreamGroups::span(g, e->grp.n, st.S, &s0, &n);
            cudaStream_t gs = e->grp.streams[g];
            BM_CUDA(cudaStreamWaitEvent(gs, e->grp.fork, 0));
            BM_CUDA(bm::botsort_step_launch(bm::byte_substate(st, s0, n, (int)E), e->pr, d_dets + s0 * D * 6, d_ndets + s0,
                                            d_embs ? d_embs + s0 * D * E : nullptr,
                                            d_warps ? d_warps + (size_t)s0 * 6 : nullptr, d_out + s0 * orow,
                                            d_nout + s0, gs));
            BM_CUDA(cudaEventRecord(e->grp.done[g], gs));
            BM_CUDA(cudaStreamWaitEvent(main, e->grp.done[g], 0));
        }
        return 0;
    }
    BM_CUDA(bm::botsort_step_launch(e->st, e->pr, d_dets, d_ndets, d_embs, d_warps, d_out, d_nout, main));
    return 0;
}

int bm_botsort_set_groups(void* handle, int groups) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || !e->bot) return fail(-1, "bm_botsort_set_groups: bad argument");
    if (e->grp.set(groups, e->st.S) != 0) return fail(-2, "bm_botsort_set_groups: stream / event creation failed");
    return e->grp.n;
}

int bm_botsort_destroy(void* handle) { return bm_bytetrack_destroy(handle); }
int bm_botsort_reset(void* handle, void* stream) { return bm_bytetrack_reset(handle, stream); }
int bm_botsort_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream) {
    return bm_bytetrack_counters(handle, err_bits, out4, stream);
}
int bm_botsort_peek(void* handle, int si, int* n_active, int* n_lost, double* h_mean, double* h_cov, int* h_ids,
                    void* stream) {
    return bm_bytetrack_peek(handle, si, n_active, n_lost, h_mean, h_cov, h_ids, stream);
}

int bm_bytetrack_destroy(void* handle) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e) return 0;
    e->grp.clear();
    for (void* p : e->allocs) cudaFree(p);
    delete e;
    return 0;
}

int bm_bytetrack_reset(void* handle, void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e) return fail(-1, "null handle");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::ByteState& st = e->st;
    const size_t S = st.S, T = st.Tcap;
    BM_CUDA(cudaMemsetAsync(st.hdr, 0, S * bm::BH_COUNT * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.n_updates, 0, S * sizeof(unsigned long long), s));
    BM_CUDA(cudaMemsetAsync(st.ints, 0, S * bm::BI_COUNT * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.act, 0, S * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.lost, 0, S * T * sizeof(int), s));
    BM_CUDA(cudaMemsetAsync(st.xP, 0, S * bm::XA_N * T * sizeof(double), s));
    BM_CUDA(cudaMemsetAsync(st.tbox, 0, S * 4 * T * sizeof(double), s));
    if (st.pc_act) {
        BM_CUDA(cudaMemsetAsync(st.pc_n, 0, S * st.pc_C * sizeof(int), s));
        BM_CUDA(cudaMemsetAsync(st.pc_meta, 0xFF, S * 2 * sizeof(int), s));
    }
    return 0;
}

int bm_bytetrack_select_class(void* handle, int n_classes, int cls, int new_frame, void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || n_classes < 1 || cls < 0 || cls >= n_classes) return fail(-1, "bm_bytetrack_select_class: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    bm::ByteState& st = e->st;
    if (!st.pc_act) {
        const size_t S = st.S;
        int *a = nullptr, *n = nullptr, *m = nullptr;
        BM_CUDA(dev_alloc(e, &a, S * n_classes * st.Tcap));
        BM_CUDA(dev_alloc(e, &n, S * n_classes));
        BM_CUDA(dev_alloc(e, &m, S * 2));
        BM_CUDA(cudaMemsetAsync(n, 0, S * n_classes * sizeof(int), s));
        BM_CUDA(cudaMemsetAsync(m, 0xFF, S * 2 * sizeof(int), s));
        st.pc_act = a; st.pc_n = n; st.pc_meta = m; st.pc_C = n_classes;
    }
    if (n_classes != st.pc_C) return fail(-1, "bm_bytetrack_select_class: the number of classes is fixed by the first call");
    byte_select_class_kernel<<<st.S, 128, 0, s>>>(st, st.pc_act, st.pc_n, st.pc_meta, st.pc_C, cls, new_frame);
    BM_CUDA(cudaGetLastError());
    return 0;
}

int bm_bytetrack_step(void* handle, const float* d_dets, const int* d_ndets, double* d_out, int* d_nout,
                      void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || e->bot || !d_dets || !d_ndets || !d_out || !d_nout) return fail(-1, "bm_bytetrack_step: bad argument");
    BM_CUDA(bm::bytetrack_step_launch(e->st, e->pr, d_dets, d_ndets, d_out, d_nout,
                                      static_cast<cudaStream_t>(stream)));
    return 0;
}

int bm_bytetrack_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || !out4) return fail(-1, "bm_bytetrack_counters: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int S = e->st.S;
    std::vector<int> hdr((size_t)S * bm::BH_COUNT);
    std::vector<unsigned long long> nu(S);
    BM_CUDA(cudaMemcpyAsync(hdr.data(), e->st.hdr, hdr.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(nu.data(), e->st.n_updates, nu.size() * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    int err = 0;
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
    for (int i = 0; i < S; i++) {
        const int* h = &hdr[(size_t)i * bm::BH_COUNT];
        err |= h[bm::BH_ERR];
        out4[0] += nu[i];
        out4[1] += (unsigned long long)h[bm::BH_LAPS];
        out4[2] += (unsigned long long)h[bm::BH_MATCH];
        out4[3] += (unsigned long long)h[bm::BH_BIRTHS];
    }
    if (err_bits) *err_bits = err;
    return 0;
}

int bm_bytetrack_peek(void* handle, int si, int* n_active, int* n_lost, double* h_mean, double* h_cov,
                      int* h_ids, void* stream) {
    ByteEngine* e = static_cast<ByteEngine*>(handle);
    if (!e || si < 0 || si >= e->st.S) return fail(-1, "bm_bytetrack_peek: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bm::ByteState& st = e->st;
    const size_t T = st.Tcap;
    std::vector<double> xp(bm::XA_N * T), dp(st.dP ? 64 * T : 0);
    std::vector<int> act(T), lost(T), ints(bm::BI_COUNT * T), hdr(bm::BH_COUNT);
    BM_CUDA(cudaMemcpyAsync(xp.data(), st.xP + (size_t)si * bm::XA_N * T, xp.size() * 8, cudaMemcpyDeviceToHost, s));
    if (st.dP) BM_CUDA(cudaMemcpyAsync(dp.data(), st.dP + (size_t)si * 64 * T, dp.size() * 8, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(act.data(), st.act + (size_t)si * T, T * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(lost.data(), st.lost + (size_t)si * T, T * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(ints.data(), st.ints + (size_t)si * bm::BI_COUNT * T, ints.size() * 4,
                            cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaMemcpyAsync(hdr.data(), st.hdr + (size_t)si * bm::BH_COUNT, hdr.size() * 4, cudaMemcpyDeviceToHost, s));
    BM_CUDA(cudaStreamSynchronize(s));
    const int na = hdr[bm::BH_NACT], nl = hdr[bm::BH_NLOST];
    if (n_active) *n_active = na;
    if (n_lost) *n_lost = nl;
    for (int p = 0; p < na + nl; p++) {
        const int slot = p < na ? act[p] : lost[p - na];
        if (h_mean) for (int c = 0; c < 8; c++) h_mean[p * 8 + c] = xp[c * T + slot];
        if (h_cov && st.dP && ints[bm::BI_DENSE * T + slot]) {
            for (int c = 0; c < 64; c++) h_cov[(size_t)p * 64 + c] = dp[c * T + slot];
        } else if (h_cov) {
            double pb[bm::XA_P];
            for (int c = 0; c < bm::XA_P; c++) pb[c] = xp[(8 + c) * T + slot];
            bm::xyah_expand(pb, h_cov + (size_t)p * 64);
        }
        if (h_ids) h_ids[p] = ints[bm::BI_ID * T + slot];
    }
    return 0;
}

}  // extern "C"
