// laser_slam_b200/csrc/zhsm_group.cu -- zhsm_group_*: the multi-GPU entry a C++ host can call.
//
// The reference has no multi-GPU code; north_star asks for batches of scan pairs "sharded across the 8 GPUs of one box,
// with NCCL over NVLink used only to all-gather result poses and covariances for the pose graph", and the host side is
// C++.  One PROCESS drives n GPUs here (ncclCommInitAll): a handle per device, jobs cut into contiguous blocks (GPU k
// owns jobs [k*B/n, (k+1)*B/n): laser_slam_b200/shard.py's rule), each GPU receives only the scans its block
// references, runs zhsm_plicp_batch on its own stream, and ONE ncclAllGather of the 128-byte records leaves the whole
// batch's records on every GPU (zhsm_group_gathered) and, from GPU 0, in the caller's host array.  No other collective,
// no peer copies of scans.  The torchrun path of bench.py (one process per GPU) is the same partition + the same gather.
//
// libnccl.so.2 is loaded with dlopen at group creation: libzhsm.so has no load-time NCCL dependency, and inside a
// process that already holds an NCCL (torch's bundled one) the same library is reused.
#include <dlfcn.h>
#include <nccl.h>  // types only: every function goes through dlsym
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "zhsm_internal.h"

namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string &err) {
        if (lib) return true;
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) { err = std::string("libnccl.so.2 not found: ") + dlerror(); return false; }
#define ZH_SYM(field, name) *(void **)(&field) = dlsym(lib, name); if (!field) { err = "libnccl: symbol " name " missing"; return false; }
        ZH_SYM(CommInitAll, "ncclCommInitAll") ZH_SYM(CommDestroy, "ncclCommDestroy") ZH_SYM(GroupStart, "ncclGroupStart")
        ZH_SYM(GroupEnd, "ncclGroupEnd") ZH_SYM(AllGather, "ncclAllGather") ZH_SYM(GetErrorString, "ncclGetErrorString")
#undef ZH_SYM
        return true;
    }
};
NcclApi g_nccl;

}  // namespace

struct zhsm_group {
    int n = 0;
    std::vector<int> devices;
    std::vector<zhsm_handle *> h;
    std::vector<ncclComm_t> comm;
    std::vector<zhsm_result *> gathered;  // per device: [n * block] records of the last batch (device memory, handle-owned)
    int last_block = 0, last_B = 0;
    std::string err;
};

extern "C" zhsm_status zhsm_group_create(const int *devices, int n, zhsm_group **out) {
    if (!out) return ZHSM_ERR_ARG;
    *out = nullptr;
    if (!devices || n < 1 || n > 64) return ZHSM_ERR_ARG;
    zhsm_group *g = new zhsm_group();
    g->n = n;
    g->devices.assign(devices, devices + n);
    g->h.assign(n, nullptr);
    g->gathered.assign(n, nullptr);
    for (int k = 0; k < n; k++) {
        zhsm_status s = zhsm_create(devices[k], &g->h[k]);
        if (s != ZHSM_OK) { zhsm_group_destroy(g); return s; }
    }
    if (!g_nccl.load(g->err)) { fprintf(stderr, "zhsm_group_create: %s\n", g->err.c_str()); zhsm_group_destroy(g); return ZHSM_ERR_UNSUPPORTED; }
    g->comm.assign(n, nullptr);
    ncclResult_t r = g_nccl.CommInitAll(g->comm.data(), n, devices);
    if (r != ncclSuccess) {
        fprintf(stderr, "zhsm_group_create: ncclCommInitAll: %s\n", g_nccl.GetErrorString(r));
        g->comm.clear();
        zhsm_group_destroy(g);
        return ZHSM_ERR_CUDA;
    }
    *out = g;
    return ZHSM_OK;
}

extern "C" void zhsm_group_destroy(zhsm_group *g) {
    if (!g) return;
    for (size_t k = 0; k < g->comm.size(); k++)
        if (g->comm[k]) { cudaSetDevice(g->devices[k]); g_nccl.CommDestroy(g->comm[k]); }
    for (int k = 0; k < g->n; k++) if (g->h[k]) zhsm_destroy(g->h[k]);
    delete g;
}

extern "C" int zhsm_group_size(zhsm_group *g) { return g ? g->n : 0; }
extern "C" zhsm_handle *zhsm_group_handle(zhsm_group *g, int k) { return (g && k >= 0 && k < g->n) ? g->h[k] : nullptr; }
extern "C" const char *zhsm_group_last_error(zhsm_group *g) { return g ? g->err.c_str() : "null group"; }

extern "C" zhsm_status zhsm_group_set_laser(zhsm_group *g, const zhsm_laser *laser) {
    if (!g || !laser) return ZHSM_ERR_ARG;
    for (int k = 0; k < g->n; k++) {
        zhsm_status s = zhsm_set_laser(g->h[k], laser);
        if (s != ZHSM_OK) { g->err = zhsm_last_error(g->h[k]); return s; }
    }
    return ZHSM_OK;
}

extern "C" const zhsm_result *zhsm_group_gathered(zhsm_group *g, int k, int *block, int *B) {
    if (!g || k < 0 || k >= g->n) return nullptr;
    if (block) *block = g->last_block;
    if (B) *B = g->last_B;
    return g->gathered[k];
}

extern "C" zhsm_status zhsm_group_plicp_batch(zhsm_group *g, const float *scans, int S, const int32_t *ref_idx,
                                              const int32_t *sens_idx, const double *first_guess, int B,
                                              const zhsm_plicp_params *params, zhsm_result *results, unsigned flags) {
    if (!g) return ZHSM_ERR_ARG;
    if (!scans || !ref_idx || !sens_idx || !results || S <= 0 || B < 0) { g->err = "zhsm_group_plicp_batch: bad argument"; return ZHSM_ERR_ARG; }
    if (flags & ~(ZHSM_FILTER_SCANS)) { g->err = "zhsm_group_plicp_batch: host arrays only; flags may hold ZHSM_FILTER_SCANS"; return ZHSM_ERR_ARG; }
    if (B == 0) return ZHSM_OK;
    for (int j = 0; j < B; j++)
        if (ref_idx[j] < 0 || ref_idx[j] >= S || sens_idx[j] < 0 || sens_idx[j] >= S) { g->err = "zhsm_group_plicp_batch: scan index out of range"; return ZHSM_ERR_ARG; }
    const int P = g->n;
    const int block = (B + P - 1) / P;  // the gather needs equal blocks: the last ones are padded (records zeroed)
    if (!g->h[0]->have_laser) { g->err = "zhsm_group_set_laser not called"; return ZHSM_ERR_ARG; }
    const int n = g->h[0]->laser.n;
    std::vector<zhsm_status> st(P, ZHSM_OK);
    std::vector<zhsm_result *> d_local(P, nullptr);
    // one host thread per GPU: the uploads of the blocks' scans run side by side, every kernel is queued without waiting
    auto work = [&](int k) {
        zhsm_handle *h = g->h[k];
        const int lo = std::min(B, k * block), hi = std::min(B, lo + block), bk = hi - lo;
        if (cudaSetDevice(h->device) != cudaSuccess) { st[k] = ZHSM_ERR_CUDA; return; }
        void *p;
        if ((st[k] = zhsm_stage(h, 23, sizeof(zhsm_result) * (size_t)block * (P + 1), &p)) != ZHSM_OK) return;
        zhsm_result *d_all = (zhsm_result *)p;        // [P * block] gathered, then [block] local
        d_local[k] = d_all + (size_t)P * block;
        g->gathered[k] = d_all;
        if (cudaMemsetAsync(d_local[k], 0, sizeof(zhsm_result) * (size_t)block, h->stream) != cudaSuccess) { st[k] = ZHSM_ERR_CUDA; return; }
        if (bk == 0) return;
        // the scans this block references: the contiguous slice [smin, smax] of the pool, indices rebased
        int smin = S, smax = -1;
        for (int j = lo; j < hi; j++) {
            smin = std::min(smin, std::min(ref_idx[j], sens_idx[j]));
            smax = std::max(smax, std::max(ref_idx[j], sens_idx[j]));
        }
        const int Sk = smax - smin + 1;
        std::vector<int32_t> idx(2 * (size_t)bk);
        for (int j = 0; j < bk; j++) { idx[j] = ref_idx[lo + j] - smin; idx[bk + j] = sens_idx[lo + j] - smin; }
        if ((st[k] = zhsm_stage(h, 20, sizeof(float) * (size_t)Sk * n, &p)) != ZHSM_OK) return;
        float *d_scans = (float *)p;
        if ((st[k] = zhsm_stage(h, 21, sizeof(int32_t) * 2 * (size_t)bk, &p)) != ZHSM_OK) return;
        int32_t *d_idx = (int32_t *)p;
        double *d_fg = nullptr;
        if (first_guess) {
            if ((st[k] = zhsm_stage(h, 22, sizeof(double) * 3 * (size_t)bk, &p)) != ZHSM_OK) return;
            d_fg = (double *)p;
        }
        bool ok = cudaMemcpyAsync(d_scans, scans + (size_t)smin * n, sizeof(float) * (size_t)Sk * n, cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
        ok = ok && cudaMemcpyAsync(d_idx, idx.data(), sizeof(int32_t) * 2 * (size_t)bk, cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
        if (d_fg) ok = ok && cudaMemcpyAsync(d_fg, first_guess + 3 * (size_t)lo, sizeof(double) * 3 * (size_t)bk, cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
        // (idx is a local vector: the copy of pageable memory is complete for the source when cudaMemcpyAsync returns)
        if (!ok) { h->err = "zhsm_group_plicp_batch: H2D copy failed"; st[k] = ZHSM_ERR_CUDA; return; }
        st[k] = zhsm_plicp_batch(h, d_scans, Sk, d_idx, d_idx + bk, d_fg, bk, params, d_local[k], nullptr, nullptr,
                                 ZHSM_ALL_ON_DEVICE | (flags & ZHSM_FILTER_SCANS), h->stream);
    };
    {
        std::vector<std::thread> th;
        for (int k = 1; k < P; k++) th.emplace_back(work, k);
        work(0);
        for (auto &t : th) t.join();
    }
    for (int k = 0; k < P; k++)
        if (st[k] != ZHSM_OK) { g->err = std::string("GPU ") + std::to_string(g->devices[k]) + ": " + zhsm_last_error(g->h[k]); return st[k]; }
    // the one collective: every GPU's block of records to every GPU, queued behind that GPU's kernel on its stream
    ncclResult_t r = g_nccl.GroupStart();
    for (int k = 0; k < P && r == ncclSuccess; k++) {
        cudaSetDevice(g->devices[k]);
        r = g_nccl.AllGather(d_local[k], g->gathered[k], sizeof(zhsm_result) * (size_t)block, ncclUint8, g->comm[k], g->h[k]->stream);
    }
    ncclResult_t r2 = g_nccl.GroupEnd();
    if (r == ncclSuccess) r = r2;
    if (r != ncclSuccess) { g->err = std::string("ncclAllGather: ") + g_nccl.GetErrorString(r); return ZHSM_ERR_CUDA; }
    // GPU 0's copy of the gathered records to the caller (the first B of P * block); then everything has drained
    cudaSetDevice(g->devices[0]);
    if (cudaMemcpyAsync(results, g->gathered[0], sizeof(zhsm_result) * (size_t)B, cudaMemcpyDeviceToHost, g->h[0]->stream) != cudaSuccess) {
        g->err = "zhsm_group_plicp_batch: D2H copy failed";
        return ZHSM_ERR_CUDA;
    }
    for (int k = 0; k < P; k++) {
        cudaSetDevice(g->devices[k]);
        if (cudaStreamSynchronize(g->h[k]->stream) != cudaSuccess) { g->err = "zhsm_group_plicp_batch: a GPU failed"; return ZHSM_ERR_CUDA; }
    }
    g->last_block = block;
    g->last_B = B;
    return ZHSM_OK;
}
