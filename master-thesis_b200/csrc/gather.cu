// gather.cu -- the one exchange step of the path: filtered clouds of every rank -> the mapping rank.
// Replaces, for frames sharded over one process per GPU, the hand-over MappingSystem::AddFrame -> MapDataBase::InsertBlock
// (cv_reconstruct/src/system/MappingSystem.cpp:344-345, MapDataBase.cpp:28-54).
//
// NCCL is bound at run time: libnccl.so.2 is looked up with dlopen (RTLD_NOLOAD first, so a process that already carries
// one -- e.g. the copy torch ships -- keeps using it).  The payload leaves from the packed device buffer the filter stages
// wrote (k_pack_clouds, api.cu); no staging copy on the sending side.
#include "cvr_internal.cuh"
#include <dlfcn.h>
#include <mutex>
#include <string.h>

// the slice of NCCL's C API used here (stable since NCCL 2.7)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[CVR_COMM_ID_BYTES]; } ncclUniqueId_;
enum { ncclSuccess_ = 0 };
enum { ncclInt8_ = 0, ncclUint8_ = 1, ncclInt32_ = 2 };
struct NcclApi {
    void* h;
    int (*GetUniqueId)(ncclUniqueId_*);
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_, int);
    int (*CommDestroy)(ncclComm_t);
    int (*CommCount)(const ncclComm_t, int*);
    int (*CommUserRank)(const ncclComm_t, int*);
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t);
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*GroupStart)(void);
    int (*GroupEnd)(void);
    const char* (*GetErrorString)(int);
};
static NcclApi g_nccl;
static std::once_flag g_nccl_once;
static std::string g_nccl_err;

static void nccl_bind()
{
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { g_nccl_err = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : ""); return; }
    NcclApi a; memset(&a, 0, sizeof(a)); a.h = h;
#define BIND(field, sym) do { *(void**)(&a.field) = dlsym(h, sym); if (!a.field) { g_nccl_err = std::string("libnccl: missing symbol ") + sym; return; } } while (0)
    BIND(GetUniqueId, "ncclGetUniqueId"); BIND(CommInitRank, "ncclCommInitRank"); BIND(CommDestroy, "ncclCommDestroy");
    BIND(CommCount, "ncclCommCount"); BIND(CommUserRank, "ncclCommUserRank"); BIND(AllGather, "ncclAllGather");
    BIND(Send, "ncclSend"); BIND(Recv, "ncclRecv"); BIND(GroupStart, "ncclGroupStart"); BIND(GroupEnd, "ncclGroupEnd");
    BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    g_nccl = a;
}
static const NcclApi* nccl_api()
{
    std::call_once(g_nccl_once, nccl_bind);
    return g_nccl.h ? &g_nccl : nullptr;
}
#define CVR_NCCL(ctx, api, call)                                                                  \
    do {                                                                                          \
        const int r__ = (call);                                                                   \
        if (r__ != ncclSuccess_) { (ctx)->err = std::string(#call) + ": " + (api)->GetErrorString(r__); return CVR_ERR_CUDA; } \
    } while (0)

void cvr_comm_reset(cvr_ctx* c)
{
    c->comm = nullptr; c->comm_owned = false; c->comm_nranks = 1; c->comm_rank = 0; c->comm_stream = nullptr;
    c->d_hdr_all = nullptr; c->h_hdr_all = nullptr; c->d_gather = nullptr; c->gather_cap = 0;
}
void cvr_comm_release(cvr_ctx* c)
{
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    if (c->comm && c->comm_owned) { const NcclApi* a = nccl_api(); if (a) a->CommDestroy((ncclComm_t)c->comm); }
    if (c->d_hdr_all) cudaFree(c->d_hdr_all);
    if (c->h_hdr_all) cudaFreeHost(c->h_hdr_all);
    if (c->d_gather) cudaFree(c->d_gather);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    cvr_comm_reset(c);
}

extern "C" int cvr_comm_get_unique_id(void* id)
{
    if (!id) return CVR_ERR_INVALID_ARG;
    const NcclApi* a = nccl_api();
    if (!a) { cvr_set_thread_error(g_nccl_err.c_str()); return CVR_ERR_UNSUPPORTED; }
    ncclUniqueId_ u;
    const int r = a->GetUniqueId(&u);
    if (r != ncclSuccess_) { cvr_set_thread_error((std::string("ncclGetUniqueId: ") + a->GetErrorString(r)).c_str()); return CVR_ERR_CUDA; }
    memcpy(id, &u, sizeof(u));
    return CVR_OK;
}

static int comm_staging(cvr_ctx* c)
{
    const size_t hdr = 2 + (size_t)c->maxB;
    if (!c->comm_stream) CVR_CUDA(c, cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
    if (c->d_hdr_all) { cudaFree(c->d_hdr_all); c->d_hdr_all = nullptr; }
    if (c->h_hdr_all) { cudaFreeHost(c->h_hdr_all); c->h_hdr_all = nullptr; }
    CVR_CUDA(c, cudaMalloc((void**)&c->d_hdr_all, (size_t)c->comm_nranks * hdr * sizeof(int)));
    CVR_CUDA(c, cudaHostAlloc((void**)&c->h_hdr_all, (size_t)c->comm_nranks * hdr * sizeof(int), cudaHostAllocDefault));
    return CVR_OK;
}

extern "C" int cvr_comm_init(cvr_ctx* c, int nranks, int rank, const void* id)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) { c->err = "cvr_comm_init: bad rank / id"; return CVR_ERR_INVALID_ARG; }
    if (c->comm) { c->err = "a communicator is already bound to this context"; return CVR_ERR_INVALID_ARG; }
    const NcclApi* a = nccl_api();
    if (!a) { c->err = g_nccl_err; return CVR_ERR_UNSUPPORTED; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    ncclUniqueId_ u; memcpy(&u, id, sizeof(u));
    ncclComm_t comm = nullptr;
    CVR_NCCL(c, a, a->CommInitRank(&comm, nranks, u, rank));
    c->comm = comm; c->comm_owned = true; c->comm_nranks = nranks; c->comm_rank = rank;
    return comm_staging(c);
}

extern "C" int cvr_comm_attach(cvr_ctx* c, void* nccl_comm)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (!nccl_comm) { c->err = "cvr_comm_attach: null communicator"; return CVR_ERR_INVALID_ARG; }
    if (c->comm) { c->err = "a communicator is already bound to this context"; return CVR_ERR_INVALID_ARG; }
    const NcclApi* a = nccl_api();
    if (!a) { c->err = g_nccl_err; return CVR_ERR_UNSUPPORTED; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    int n = 0, r = 0;
    CVR_NCCL(c, a, a->CommCount((ncclComm_t)nccl_comm, &n));
    CVR_NCCL(c, a, a->CommUserRank((ncclComm_t)nccl_comm, &r));
    c->comm = nccl_comm; c->comm_owned = false; c->comm_nranks = n; c->comm_rank = r;
    return comm_staging(c);
}

extern "C" int cvr_comm_destroy(cvr_ctx* c)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    cudaSetDevice(c->device);
    cvr_comm_release(c);
    return CVR_OK;
}

// Host arithmetic of the gather (no GPU involved): where each rank's points and frames land on the root.
extern "C" int cvr_gather_layout(const int32_t* headers, int nranks, int header_ints, int64_t* rank_offset, int32_t* frame_offsets,
                                 int frame_offsets_capacity, int32_t* frames_of_rank)
{
    if (!headers || nranks < 1 || header_ints < 2 || !rank_offset) return CVR_ERR_INVALID_ARG;
    int64_t pts = 0; int frames = 0;
    for (int r = 0; r < nranks; r++) {
        const int32_t* h = headers + (size_t)r * header_ints;
        const int nf = h[0];
        if (nf < 0 || nf > header_ints - 2 || h[1] < 0) return CVR_ERR_INVALID_ARG;
        int64_t sum = 0;
        for (int i = 0; i < nf; i++) {
            if (h[2 + i] < 0) return CVR_ERR_INVALID_ARG;
            if (frame_offsets) {
                if (frames + i >= frame_offsets_capacity) return CVR_ERR_CAPACITY;
                if (pts + sum > 0x7fffffffLL) return CVR_ERR_CAPACITY;
                frame_offsets[frames + i] = (int32_t)(pts + sum);
            }
            sum += h[2 + i];
        }
        if (sum != h[1]) return CVR_ERR_INVALID_ARG;                   // a header whose total disagrees with its counts is corrupt
        rank_offset[r] = pts;
        if (frames_of_rank) frames_of_rank[r] = nf;
        pts += sum; frames += nf;
    }
    rank_offset[nranks] = pts;
    if (frame_offsets) {
        if (frames >= frame_offsets_capacity || pts > 0x7fffffffLL) return CVR_ERR_CAPACITY;
        frame_offsets[frames] = (int32_t)pts;
    }
    return frames;
}

extern "C" int cvr_gather_clouds(cvr_ctx* c, int root, int mem, cvr_point* out, size_t out_capacity, int32_t* frame_offsets, int frame_offsets_capacity,
                                 int32_t* frames_of_rank)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (!c->lp_valid) { c->err = "cvr_gather_clouds: no completed batch with a cloud stage to gather"; return CVR_ERR_INVALID_ARG; }
    const int nranks = c->comm ? c->comm_nranks : 1, rank = c->comm ? c->comm_rank : 0;
    if (root < 0 || root >= nranks) { c->err = "cvr_gather_clouds: root out of range"; return CVR_ERR_INVALID_ARG; }
    const bool isRoot = rank == root;
    if (isRoot && !out) { c->err = "cvr_gather_clouds: the root needs an output buffer"; return CVR_ERR_INVALID_ARG; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    const int hdr = 2 + c->maxB;
    const NcclApi* a = nullptr;
    if (!c->comm) {                                                    // one rank: the header is the local one
        if (!c->comm_stream) { c->comm_nranks = 1; c->comm_rank = 0; const int rc0 = comm_staging(c); if (rc0) return rc0; }
        CVR_CUDA(c, cudaMemcpyAsync(c->h_hdr_all, c->lp_dhdr, (size_t)hdr * sizeof(int), cudaMemcpyDeviceToHost, c->comm_stream));
    } else {
        a = nccl_api();
        if (!a) { c->err = g_nccl_err; return CVR_ERR_UNSUPPORTED; }
        CVR_NCCL(c, a, a->AllGather(c->lp_dhdr, c->d_hdr_all, (size_t)hdr, ncclInt32_, (ncclComm_t)c->comm, c->comm_stream));
        CVR_CUDA(c, cudaMemcpyAsync(c->h_hdr_all, c->d_hdr_all, (size_t)nranks * hdr * sizeof(int), cudaMemcpyDeviceToHost, c->comm_stream));
    }
    CVR_CUDA(c, cudaStreamSynchronize(c->comm_stream));                // the receiver has to know the sizes before it can post receives
    std::vector<int64_t> roff((size_t)nranks + 1);
    const int frames = cvr_gather_layout(c->h_hdr_all, nranks, hdr, roff.data(), isRoot ? frame_offsets : nullptr, frame_offsets_capacity,
                                         isRoot ? frames_of_rank : nullptr);
    if (frames < 0) {
        // every rank sees the same headers, so every rank fails here alike when a header is corrupt; a too-small frame_offsets
        // only fails the root, which must still take part in the exchange below -> report it afterwards
        if (!(isRoot && frames == CVR_ERR_CAPACITY)) { c->err = "cvr_gather_clouds: inconsistent headers (do all ranks use the same max_batch?)"; return frames; }
        cvr_gather_layout(c->h_hdr_all, nranks, hdr, roff.data(), nullptr, 0, nullptr);
    }
    const size_t total = (size_t)roff[nranks];
    int status = frames < 0 ? frames : CVR_OK;
    cvr_point* dst = nullptr;
    if (isRoot) {
        if (total > out_capacity) status = CVR_ERR_CAPACITY;
        if (mem == CVR_MEM_DEVICE && status == CVR_OK) dst = out;
        else {                                                         // host output (or an undersized buffer): receive into library memory
            if (c->gather_cap < total) {
                if (c->d_gather) { cudaFree(c->d_gather); c->d_gather = nullptr; c->gather_cap = 0; }
                const size_t want = total + total / 2 + 1024;
                CVR_CUDA(c, cudaMalloc((void**)&c->d_gather, want * sizeof(cvr_point)));
                c->gather_cap = want;
            }
            dst = c->d_gather;
        }
    }
    const size_t mine = (size_t)(roff[rank + 1] - roff[rank]);
    if (a && nranks > 1) {
        CVR_NCCL(c, a, a->GroupStart());
        if (isRoot) {
            for (int r = 0; r < nranks; r++) {
                const size_t n = (size_t)(roff[r + 1] - roff[r]);
                if (r != root && n) CVR_NCCL(c, a, a->Recv(dst + roff[r], n * sizeof(cvr_point), ncclUint8_, r, (ncclComm_t)c->comm, c->comm_stream));
            }
        } else if (mine) {
            CVR_NCCL(c, a, a->Send(c->lp_pts, mine * sizeof(cvr_point), ncclUint8_, root, (ncclComm_t)c->comm, c->comm_stream));
        }
        CVR_NCCL(c, a, a->GroupEnd());
    }
    if (isRoot) {
        if (mine) CVR_CUDA(c, cudaMemcpyAsync(dst + roff[rank], c->lp_pts, mine * sizeof(cvr_point), cudaMemcpyDeviceToDevice, c->comm_stream));
        if (status == CVR_OK && mem != CVR_MEM_DEVICE && total)
            CVR_CUDA(c, cudaMemcpyAsync(out, dst, total * sizeof(cvr_point), cudaMemcpyDeviceToHost, c->comm_stream));
    }
    CVR_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    if (status == CVR_ERR_CAPACITY) c->err = "cvr_gather_clouds: out_capacity / frame_offsets_capacity too small for the gathered clouds";
    return status;
}
