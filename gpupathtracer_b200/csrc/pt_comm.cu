// pt_comm.cu -- the sample split across GPUs behind the C ABI.
//
// The path shards by SAMPLES: scene + BVH replicated, every rank renders its share of the samples of every pixel with its own
// RNG table into a private RGBA32F sum, and ONE ncclReduce per batch brings the sums to rank 0 (Iterative.comp:263,295
// accumulate a sum, Present.frag:30 divides by the sample count -- sums over ranks add the same way). No other exchange on
// the data path. Two ways in:
//   pt_comm_init     one process per GPU (torchrun / MPI): the caller moves a 128-byte NCCL id from rank 0 to the others
//   pt_create_multi  one process, one host thread driving n GPUs (ncclCommInitAll), the seam b200pt::Renderer uses
// A batch renders into one of two accumulators; its reduce runs on a dedicated stream and overlaps the next batch, which
// renders into the other accumulator. Rank 0 keeps the running total (total += reduced batch).
// NCCL is bound with dlopen (the copy already in the process -- e.g. torch's -- if there is one), so the library has no
// link-time dependency on it and single-GPU users never load it.
#include "pt_context.h"
#include <nccl.h>
#include <dlfcn.h>
#include <cstring>
#include <cstdlib>

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    void* h = nullptr;
    if (const char* e = getenv("PT_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);         // a copy the process already holds (torch's)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return nullptr;
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) ok = false; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
    api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
    if (!ok) return nullptr;
    api.handle = h;
    return &api;
}

#define NK(call)                                                                                                   \
    do {                                                                                                           \
        ncclResult_t r__ = (call);                                                                                 \
        if (r__ != ncclSuccess) return fail(PT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, nccl_api()->GetErrorString(r__), __FILE__, __LINE__); \
    } while (0)

__global__ void k_add_into(float4* __restrict__ total, const float4* __restrict__ batch, long long n) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 t = total[i], b = batch[i];
    total[i] = make_float4(t.x + b.x, t.y + b.y, t.z + b.z, 1.0f);      // alpha stays 1 (Iterative.comp:263)
}

}  // namespace

struct pt_comm_state {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t stream = nullptr;                 // the reduce (+ rank 0's add) runs here, beside the next batch
    DevBuf<float4> accum1;                         // second batch accumulator (the first is the context's own)
    DevBuf<float4> reduced, total;                 // rank 0: sum over ranks of one batch; running total
    cudaEvent_t ev_rendered[2] = {nullptr, nullptr}, ev_reduced[2] = {nullptr, nullptr};
    cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;  // timing of the last reduce
    cudaEvent_t ev_done = nullptr;                 // end of the last batch on the communication stream (reduce + rank 0's add)
    int turn = 0, last_turn = -1;
    uint32_t total_samples = 0;                    // samples per pixel in `total`: world x frames of every batch
    uint64_t batches = 0;
};

// ---------------------------------------------------------------------------------------------- per-context communicator
static float4* batch_accum(pt_context* c, int k) { return k == 0 ? c->d_accum_own.p : c->comm->accum1.p; }

static int comm_attach(pt_context* c, ncclComm_t comm, int rank, int world) {
    CK(cudaSetDevice(c->cfg.device));
    pt_comm_state* s = new pt_comm_state();
    s->comm = comm; s->rank = rank; s->world = world;
    c->comm = s;
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CK(s->accum1.alloc(npx));
    if (rank == 0) {
        CK(s->reduced.alloc(npx));
        CK(s->total.alloc(npx));
        CK(cudaMemsetAsync(s->total.p, 0, npx * sizeof(float4), c->stream));
    }
    for (int k = 0; k < 2; k++) {
        CK(cudaEventCreateWithFlags(&s->ev_rendered[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&s->ev_reduced[k], cudaEventDisableTiming));
    }
    CK(cudaEventCreate(&s->ev_r0));
    CK(cudaEventCreate(&s->ev_r1));
    CK(cudaEventCreate(&s->ev_done));
    pt_invalidate_graphs(c);                       // batches switch accumulators: frames are enqueued eagerly from here on
    c->d_accum = c->d_accum_own.p;
    CK(cudaStreamSynchronize(c->stream));
    return PT_OK;
}

void pt_comm_release(pt_context* c) {
    if (!c || !c->comm) return;
    pt_comm_state* s = c->comm;
    cudaSetDevice(c->cfg.device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->comm && nccl_api()) nccl_api()->CommDestroy(s->comm);
    s->accum1.release(); s->reduced.release(); s->total.release();
    for (int k = 0; k < 2; k++) { if (s->ev_rendered[k]) cudaEventDestroy(s->ev_rendered[k]); if (s->ev_reduced[k]) cudaEventDestroy(s->ev_reduced[k]); }
    if (s->ev_r0) cudaEventDestroy(s->ev_r0);
    if (s->ev_r1) cudaEventDestroy(s->ev_r1);
    if (s->ev_done) cudaEventDestroy(s->ev_done);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
    c->comm = nullptr;
    c->d_accum = c->d_accum_own.p;
}

// phase 1 of a batch: n frames into this rank's current batch accumulator (zeroed first), on the render streams
// Chained batches: every slice clears and renders its own band of the batch's accumulator behind its previous frame, nothing is
// joined into the main stream, and the reduce below takes the bands one by one as they finish -- no drain of the GPU between
// batches (what per-slice present is to a single GPU). Needs banded slices; stage timing keeps the joined form.
static bool batch_chained(pt_context* c) { return c->lazy_join && c->slices_banded && !c->stage_timing; }
static int batch_render(pt_context* c, int n_frames) {
    pt_comm_state* s = c->comm;
    CK(cudaSetDevice(c->cfg.device));
    int rc = pt_check_scene(c, "pt_render_batch", true);
    if (rc != PT_OK) return rc;
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int k = s->turn;
    float4* A = batch_accum(c, k);
    if (batch_chained(c)) {
        if (c->slices_presenting && (rc = pt_join_slices(c)) != PT_OK) return rc;      // a single-GPU present is still resolving
        if (!c->pending_join) CK(cudaEventRecord(c->ev0, c->stream));
        c->d_accum = A;
        // ev_reduced[k]: the reduce that last read this accumulator (two batches ago)
        c->counters.kernel_launches += pt_enqueue_frames_chained(c, n_frames, s->ev_reduced[k], true);
    } else {
        if ((rc = pt_join_slices(c)) != PT_OK) return rc;
        CK(cudaEventRecord(c->ev0, c->stream));
        CK(cudaStreamWaitEvent(c->stream, s->ev_reduced[k], 0));
        CK(cudaMemsetAsync(A, 0, npx * sizeof(float4), c->stream));
        c->d_accum = A;
        c->counters.kernel_launches += pt_enqueue_frames(c, n_frames);
        CK(cudaEventRecord(c->ev1, c->stream));
        CK(cudaEventRecord(s->ev_rendered[k], c->stream));
    }
    CK(cudaGetLastError());
    c->num_samples += n_frames;
    c->counters.frames += n_frames;
    return PT_OK;
}
// phase 2: the batch's sum to rank 0 on the communication stream (inside a group when one thread drives several ranks)
static int batch_reduce(pt_context* c, int n_frames) {
    pt_comm_state* s = c->comm;
    NcclApi* n = nccl_api();
    CK(cudaSetDevice(c->cfg.device));
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int k = s->turn;
    float4* A = batch_accum(c, k);
    if (batch_chained(c)) {
        // one reduce per band, each behind its own slice only (every rank issues them in slice order). Inside an NCCL group
        // (one thread driving several ranks) the waits all land before the grouped launch: still correct, just not staggered.
        for (int sl = 0; sl < c->n_slices; sl++) {
            const size_t first = (size_t)sl * c->slice_size;
            const size_t count = npx - first < (size_t)c->slice_size ? npx - first : (size_t)c->slice_size;
            CK(cudaStreamWaitEvent(s->stream, c->slice_done[2 * sl + 1], 0));
            if (sl == 0) CK(cudaEventRecord(s->ev_r0, s->stream));
            NK(n->Reduce(A + first, s->rank == 0 ? s->reduced.p + first : nullptr, count * 4, ncclFloat32, ncclSum, 0, s->comm, s->stream));
        }
        return PT_OK;
    }
    CK(cudaStreamWaitEvent(s->stream, s->ev_rendered[k], 0));
    CK(cudaEventRecord(s->ev_r0, s->stream));
    NK(n->Reduce(A, s->rank == 0 ? s->reduced.p : nullptr, npx * 4, ncclFloat32, ncclSum, 0, s->comm, s->stream));
    return PT_OK;
}
// phase 3: rank 0 folds the batch into the running total; everybody marks the accumulator free
static int batch_finish(pt_context* c, int n_frames) {
    pt_comm_state* s = c->comm;
    CK(cudaSetDevice(c->cfg.device));
    const long long npx = (long long)c->cfg.width * c->cfg.height;
    const int k = s->turn;
    CK(cudaEventRecord(s->ev_r1, s->stream));
    if (s->rank == 0) {
        k_add_into<<<(unsigned)((npx + 255) / 256), 256, 0, s->stream>>>(s->total.p, s->reduced.p, npx);
        c->counters.kernel_launches++;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(s->ev_reduced[k], s->stream));
    CK(cudaEventRecord(s->ev_done, s->stream));
    s->last_turn = k;
    s->turn ^= 1;
    s->total_samples += (uint32_t)n_frames * (uint32_t)s->world;
    s->batches++;
    return PT_OK;
}

static int need_comm(pt_context* c, const char* who) {
    if (!c) return fail(PT_ERR_INVALID, "%s: null context", who);
    if (!c->comm) return fail(PT_ERR_INVALID, "%s: no communicator (pt_comm_init / pt_create_multi first)", who);
    return PT_OK;
}

extern "C" {

uint64_t pt_rank_seed(uint64_t seed, int32_t rank) {
    // splitmix64 of (seed, rank): decorrelated RNG tables per GPU (SURVEY 8e "seed xor r"); rank 0 of a 1-GPU run keeps `seed`
    if (rank == 0) return seed;
    uint64_t x = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(rank + 1);
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

void pt_split_samples(uint32_t total_spp, int32_t world, int32_t rank, uint32_t* first, uint32_t* count) {
    // samples [first, first + count) of every pixel belong to `rank`: total / world each, the remainder to the lowest ranks
    if (world < 1) world = 1;
    uint32_t base = total_spp / (uint32_t)world, rem = total_spp % (uint32_t)world;
    uint32_t r = (uint32_t)(rank < 0 ? 0 : rank);
    if (count) *count = base + (r < rem ? 1u : 0u);
    if (first) *first = r * base + (r < rem ? r : rem);
}

int pt_comm_unique_id(uint8_t id[PT_COMM_ID_BYTES]) {
    if (!id) return fail(PT_ERR_INVALID, "pt_comm_unique_id: null argument");
    NcclApi* n = nccl_api();
    if (!n) return fail(PT_ERR_UNSUPPORTED, "pt_comm_unique_id: libnccl.so.2 could not be loaded (%s)", dlerror() ? dlerror() : "missing symbols");
    static_assert(sizeof(ncclUniqueId) == PT_COMM_ID_BYTES, "NCCL unique id size");
    ncclUniqueId u;
    NK(n->GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return PT_OK;
}

int pt_comm_init(pt_context* c, const uint8_t id[PT_COMM_ID_BYTES], int32_t rank, int32_t world) {
    if (!c || !id || world < 1 || rank < 0 || rank >= world) return fail(PT_ERR_INVALID, "pt_comm_init: bad argument");
    if (c->comm) return fail(PT_ERR_INVALID, "pt_comm_init: the context already has a communicator");
    NcclApi* n = nccl_api();
    if (!n) return fail(PT_ERR_UNSUPPORTED, "pt_comm_init: libnccl.so.2 could not be loaded");
    CK(cudaSetDevice(c->cfg.device));
    ncclUniqueId u; memcpy(&u, id, sizeof(u));
    ncclComm_t comm = nullptr;
    NK(n->CommInitRank(&comm, world, u, rank));
    return comm_attach(c, comm, rank, world);
}

int32_t pt_comm_rank(pt_context* c) { return c && c->comm ? c->comm->rank : 0; }
int32_t pt_comm_world(pt_context* c) { return c && c->comm ? c->comm->world : 1; }
int pt_comm_nccl_version(void) { NcclApi* n = nccl_api(); int v = 0; if (n) n->GetVersion(&v); return v; }

int pt_comm_seed_rng(pt_context* c, uint64_t seed) { return pt_seed_rng(c, pt_rank_seed(seed, pt_comm_rank(c))); }

int pt_render_batch(pt_context* c, int32_t n_frames) {
    int rc = need_comm(c, "pt_render_batch");
    if (rc != PT_OK) return rc;
    if (n_frames < 1) return fail(PT_ERR_INVALID, "pt_render_batch: n_frames must be positive");
    if ((rc = batch_render(c, n_frames)) != PT_OK) return rc;
    if ((rc = batch_reduce(c, n_frames)) != PT_OK) return rc;
    return batch_finish(c, n_frames);
}

int pt_comm_reset(pt_context* c) {
    int rc = need_comm(c, "pt_comm_reset");
    if (rc != PT_OK) return rc;
    pt_comm_state* s = c->comm;
    CK(cudaSetDevice(c->cfg.device));
    if (s->rank == 0) CK(cudaMemsetAsync(s->total.p, 0, (size_t)c->cfg.width * c->cfg.height * sizeof(float4), s->stream));
    s->total_samples = 0;
    c->num_samples = 0;
    return PT_OK;
}
uint32_t pt_comm_num_samples(pt_context* c) { return c && c->comm ? c->comm->total_samples : pt_num_samples(c); }

int pt_comm_synchronize(pt_context* c) {
    int rc = need_comm(c, "pt_comm_synchronize");
    if (rc != PT_OK) return rc;
    CK(cudaSetDevice(c->cfg.device));
    if ((rc = pt_join_slices(c)) != PT_OK) return rc;                   // chained batches: the slices' streams into the main one
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaStreamSynchronize(c->comm->stream));
    CK(cudaStreamSynchronize(c->copy_stream));
    return PT_OK;
}

int pt_comm_download_total(pt_context* c, float* rgba) {
    int rc = need_comm(c, "pt_comm_download_total");
    if (rc != PT_OK) return rc;
    if (c->comm->rank != 0) return PT_OK;                               // only rank 0 holds the total; others: nothing written
    if (!rgba) return fail(PT_ERR_INVALID, "pt_comm_download_total: null buffer on rank 0");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(rgba, c->comm->total.p, (size_t)c->cfg.width * c->cfg.height * 16, cudaMemcpyDeviceToHost, c->comm->stream));
    CK(cudaStreamSynchronize(c->comm->stream));
    return PT_OK;
}
int pt_comm_download_batch(pt_context* c, float* rgba) {
    // this rank's own sum of the LAST batch (still in its accumulator until the batch after next starts): what went into the reduce
    int rc = need_comm(c, "pt_comm_download_batch");
    if (rc != PT_OK) return rc;
    if (!rgba || c->comm->last_turn < 0) return fail(PT_ERR_INVALID, "pt_comm_download_batch: no batch rendered yet");
    CK(cudaSetDevice(c->cfg.device));
    if ((rc = pt_join_slices(c)) != PT_OK) return rc;
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpy(rgba, batch_accum(c, c->comm->last_turn), (size_t)c->cfg.width * c->cfg.height * 16, cudaMemcpyDeviceToHost));
    return PT_OK;
}
void* pt_comm_total_device_ptr(pt_context* c) { return c && c->comm && c->comm->rank == 0 ? (void*)c->comm->total.p : nullptr; }

float pt_comm_last_reduce_ms(pt_context* c) {
    if (!c || !c->comm || c->comm->batches == 0) return -1.0f;
    cudaSetDevice(c->cfg.device);
    if (cudaEventSynchronize(c->comm->ev_r1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->comm->ev_r0, c->comm->ev_r1) != cudaSuccess) return -1.0f;
    return ms;
}

// device time of the last pt_render_batch on this rank: from its first kernel to the end of its reduce (+ rank 0's add)
float pt_comm_last_batch_ms(pt_context* c) {
    if (!c || !c->comm || c->comm->batches == 0) return -1.0f;
    cudaSetDevice(c->cfg.device);
    if (cudaEventSynchronize(c->comm->ev_done) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->comm->ev_done) != cudaSuccess) return -1.0f;
    return ms;
}

// Present of the combined image (rank 0): total / total_samples through the resolve kernel on the communication stream, right
// behind the last reduce, 8-bit frame to host memory on the copy stream (double-buffered like pt_present_async)
int pt_comm_present_async(pt_context* c, uint8_t* host_rgb_u8) {
    int rc = need_comm(c, "pt_comm_present_async");
    if (rc != PT_OK) return rc;
    pt_comm_state* s = c->comm;
    if (s->rank != 0) return PT_OK;
    if (!host_rgb_u8 || s->total_samples == 0) return fail(PT_ERR_INVALID, "pt_comm_present_async: nothing rendered or null buffer");
    CK(cudaSetDevice(c->cfg.device));
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int k = c->present_turn; c->present_turn ^= 1;
    CK(c->d_present[k].alloc(npx * 3));
    CK(cudaStreamWaitEvent(s->stream, c->ev_copied[k], 0));
    launch_resolve(s->total.p, (long long)npx, s->total_samples, nullptr, c->d_present[k].p, s->stream, (c->cfg.parity_flags & PT_PARITY_LIBM) != 0);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev_resolved, s->stream));
    CK(cudaStreamWaitEvent(c->copy_stream, c->ev_resolved, 0));
    CK(cudaMemcpyAsync(host_rgb_u8, c->d_present[k].p, npx * 3, cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaEventRecord(c->ev_copied[k], c->copy_stream));
    return PT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------- BVH: build once, broadcast
// Rank `root` holds a tree (any builder); the others receive the traversal layout itself (pair records + leaf stream: usable
// the moment the broadcast lands, no host work) and the serialised form (their host mirror is fetched lazily, as after a
// device build). Triangles must already be set on every rank (the leaf stream names them by index only for shading).
static int bvh_broadcast_enqueue(pt_context* c, int root, long long* hdr_host, int phase) {
    pt_comm_state* s = c->comm;
    NcclApi* n = nccl_api();
    CK(cudaSetDevice(c->cfg.device));
    if (phase == 0) {                                                   // sizes
        CK(c->d_cursor.alloc(16));                                       // 16 ints = 8 long long
        if (s->rank == root) CK(cudaMemcpyAsync(c->d_cursor.p, hdr_host, 64, cudaMemcpyHostToDevice, c->stream));
        NK(n->Broadcast(c->d_cursor.p, c->d_cursor.p, 64, ncclUint8, root, s->comm, c->stream));
        CK(cudaMemcpyAsync(hdr_host, c->d_cursor.p, 64, cudaMemcpyDeviceToHost, c->stream));
        return PT_OK;
    }
    const size_t n_pairs = (size_t)hdr_host[0], n_leaf = (size_t)hdr_host[1], n_nodes = (size_t)hdr_host[2], n_refs = (size_t)hdr_host[3];
    NK(n->Broadcast(c->d_pairs.p, c->d_pairs.p, n_pairs * sizeof(BvhPair), ncclUint8, root, s->comm, c->stream));
    NK(n->Broadcast(c->d_leaf.p, c->d_leaf.p, n_leaf * sizeof(LeafTri), ncclUint8, root, s->comm, c->stream));
    NK(n->Broadcast(c->d_ser_nodes.p, c->d_ser_nodes.p, n_nodes * 32, ncclUint8, root, s->comm, c->stream));
    NK(n->Broadcast(c->d_ser_refs.p, c->d_ser_refs.p, n_refs * 4, ncclUint8, root, s->comm, c->stream));
    return PT_OK;
}
static int bvh_prepare_root(pt_context* c, long long hdr[8]) {
    CK(cudaSetDevice(c->cfg.device));
    if (!c->d_pairs.p || !c->d_leaf.p) return fail(PT_ERR_INVALID, "pt_comm_broadcast_bvh: the root rank has no BVH");
    if (!c->host_bvh_stale) {                                           // host-built tree: put its serialised form on the device
        const size_t nn = c->h_bvh.nodes.size() / 8, nr = c->h_bvh.refs.size();
        CK(c->d_ser_nodes.alloc(2 * nn)); CK(c->d_ser_refs.alloc(nr));
        CK(cudaMemcpyAsync(c->d_ser_nodes.p, c->h_bvh.nodes.data(), nn * 32, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_ser_refs.p, c->h_bvh.refs.data(), nr * 4, cudaMemcpyHostToDevice, c->stream));
        c->ser_nodes = (int64_t)nn; c->ser_refs = (int64_t)nr;
    }
    hdr[0] = (c->ser_nodes - 1) / 2; hdr[1] = c->n_leaf; hdr[2] = c->ser_nodes; hdr[3] = c->ser_refs;
    hdr[4] = c->h_bvh.max_depth; hdr[5] = c->h_bvh.n_leaves; hdr[6] = c->top_pairs; hdr[7] = 0;
    return PT_OK;
}
static int bvh_prepare_receiver(pt_context* c, const long long hdr[8]) {
    CK(cudaSetDevice(c->cfg.device));
    if (c->h_tris.empty()) return fail(PT_ERR_INVALID, "pt_comm_broadcast_bvh: call pt_set_triangles on every rank first");
    if (hdr[0] < 1 || hdr[1] < 1 || hdr[2] < 3 || hdr[3] < 1) return fail(PT_ERR_INVALID, "pt_comm_broadcast_bvh: the root sent an empty tree");
    pt_invalidate_graphs(c);
    CK(c->d_pairs.alloc((size_t)hdr[0])); CK(c->d_leaf.alloc((size_t)hdr[1]));
    CK(c->d_ser_nodes.alloc(2 * (size_t)hdr[2])); CK(c->d_ser_refs.alloc((size_t)hdr[3]));
    return PT_OK;
}
static int bvh_finish_receiver(pt_context* c, const long long hdr[8]) {
    c->n_leaf = (int)hdr[1]; c->ser_nodes = hdr[2]; c->ser_refs = hdr[3];
    c->h_bvh = HostBvh();
    c->h_bvh.max_depth = (int)hdr[4]; c->h_bvh.n_leaves = (int)hdr[5];
    c->top_pairs = (int)hdr[6];
    c->host_bvh_stale = true;                                           // mirror fetched when somebody asks (pt_get_bvh, wide collapse)
    return PT_OK;
}

extern "C" int pt_comm_broadcast_bvh(pt_context* c, int32_t root) {
    int rc = need_comm(c, "pt_comm_broadcast_bvh");
    if (rc != PT_OK) return rc;
    pt_comm_state* s = c->comm;
    if (root < 0 || root >= s->world) return fail(PT_ERR_INVALID, "pt_comm_broadcast_bvh: root %d of %d", root, s->world);
    if (c->cfg.bvh_layout != PT_BVH_REFERENCE) return fail(PT_ERR_UNSUPPORTED, "pt_comm_broadcast_bvh: binary layout only (the wide collapse runs per rank)");
    long long hdr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (s->rank == root && (rc = bvh_prepare_root(c, hdr)) != PT_OK) hdr[7] = -1;    // still take part: the others must not hang
    int rc2 = bvh_broadcast_enqueue(c, root, hdr, 0);
    if (rc2 != PT_OK) return rc2;
    CK(cudaStreamSynchronize(c->stream));
    if (hdr[7] != 0) return rc != PT_OK ? rc : fail(PT_ERR_INVALID, "pt_comm_broadcast_bvh: the root rank has no BVH");
    if (s->rank != root && (rc = bvh_prepare_receiver(c, hdr)) != PT_OK) return rc;
    if ((rc = bvh_broadcast_enqueue(c, root, hdr, 1)) != PT_OK) return rc;
    CK(cudaStreamSynchronize(c->stream));
    if (s->rank != root) bvh_finish_receiver(c, hdr);
    return PT_OK;
}

// ---------------------------------------------------------------------------------------------- one process, n GPUs
struct pt_group { std::vector<pt_context*> ctx; };

#define GROUP_EACH(g, expr) do { for (pt_context* c : (g)->ctx) { int rc__ = (expr); if (rc__ != PT_OK) return rc__; } } while (0)

extern "C" {

int pt_create_multi(const pt_config* cfg, const int32_t* device_ids, int32_t n, pt_group** out) {
    if (!cfg || !out || n < 1 || n > 64) return fail(PT_ERR_INVALID, "pt_create_multi: bad argument");
    *out = nullptr;
    NcclApi* api = nccl_api();
    if (n > 1 && !api) return fail(PT_ERR_UNSUPPORTED, "pt_create_multi: libnccl.so.2 could not be loaded");
    pt_group* g = new pt_group();
    std::vector<int> devs(n);
    for (int i = 0; i < n; i++) {
        devs[i] = device_ids ? device_ids[i] : i;
        pt_config one = *cfg; one.device = devs[i];
        pt_context* c = nullptr;
        int rc = pt_create(&one, &c);
        if (rc != PT_OK) { for (pt_context* d : g->ctx) pt_destroy(d); delete g; return rc; }
        g->ctx.push_back(c);
    }
    if (n > 1) {
        std::vector<ncclComm_t> comms(n, nullptr);
        ncclResult_t r = api->CommInitAll(comms.data(), n, devs.data());
        if (r != ncclSuccess) {
            for (pt_context* d : g->ctx) pt_destroy(d);
            delete g;
            return fail(PT_ERR_CUDA, "ncclCommInitAll failed: %s", api->GetErrorString(r));
        }
        for (int i = 0; i < n; i++) {
            int rc = comm_attach(g->ctx[i], comms[i], i, n);
            if (rc != PT_OK) { for (pt_context* d : g->ctx) pt_destroy(d); delete g; return rc; }
        }
    }
    *out = g;
    return PT_OK;
}

void pt_group_destroy(pt_group* g) {
    if (!g) return;
    for (pt_context* c : g->ctx) pt_destroy(c);
    delete g;
}
int32_t pt_group_size(pt_group* g) { return g ? (int32_t)g->ctx.size() : 0; }
pt_context* pt_group_context(pt_group* g, int32_t i) { return g && i >= 0 && i < (int32_t)g->ctx.size() ? g->ctx[i] : nullptr; }

// Context 0 holds the scene (pt_load_scene or the pt_set_* calls + a BVH); the others get copies: tables from the host mirrors,
// the BVH as built (never rebuilt), the environment as it is on the device.
int pt_group_share_scene(pt_group* g) {
    if (!g || g->ctx.empty()) return fail(PT_ERR_INVALID, "pt_group_share_scene: null group");
    pt_context* a = g->ctx[0];
    if (a->h_tris.empty() || a->h_mats.empty() || a->h_lights.empty()) return fail(PT_ERR_INVALID, "pt_group_share_scene: context 0 has no scene");
    int rc = pt_ensure_host_bvh(a);
    if (rc != PT_OK) return rc;
    if (a->h_bvh.nodes.empty()) return fail(PT_ERR_INVALID, "pt_group_share_scene: context 0 has no BVH");
    std::vector<pt_texture> tex(a->h_texels.size());
    for (size_t i = 0; i < tex.size(); i++) { tex[i].width = a->h_texdims[i].first; tex[i].height = a->h_texdims[i].second; tex[i].rgba = a->h_texels[i].data(); }
    std::vector<float> env;
    if (a->env_size) { env.resize(6 * (size_t)a->env_size * a->env_size * 4); if ((rc = pt_download_environment(a, env.data())) != PT_OK) return rc; }
    for (size_t i = 1; i < g->ctx.size(); i++) {
        pt_context* c = g->ctx[i];
        if ((rc = pt_set_triangles(c, a->h_tris.data(), (int64_t)(a->h_tris.size() / 20))) != PT_OK) return rc;
        if ((rc = pt_set_materials(c, a->h_mats.data(), (int32_t)a->h_mats.size(), tex.empty() ? nullptr : tex.data(), (int32_t)tex.size())) != PT_OK) return rc;
        if ((rc = pt_set_lights(c, a->h_lights.data(), (int32_t)a->h_lights.size(), a->total_light_area)) != PT_OK) return rc;
        if (a->env_size) rc = pt_set_environment_cubemap(c, env.data(), a->env_size); else rc = pt_set_environment_color(c, a->env);
        if (rc != PT_OK) return rc;
        if ((rc = pt_set_sun(c, a->sun_dir, a->sun_radius)) != PT_OK) return rc;
        if ((rc = pt_set_bvh(c, a->h_bvh.nodes.data(), (int32_t)(a->h_bvh.nodes.size() / 8), a->h_bvh.refs.data(), (int32_t)a->h_bvh.refs.size())) != PT_OK) return rc;
    }
    return PT_OK;
}

int pt_group_set_camera(pt_group* g, const pt_camera* cam) {
    if (!g) return fail(PT_ERR_INVALID, "pt_group_set_camera: null group");
    GROUP_EACH(g, pt_set_camera(c, cam));
    return PT_OK;
}
int pt_group_seed_rng(pt_group* g, uint64_t seed) {
    if (!g) return fail(PT_ERR_INVALID, "pt_group_seed_rng: null group");
    for (size_t i = 0; i < g->ctx.size(); i++) { int rc = pt_seed_rng(g->ctx[i], pt_rank_seed(seed, (int32_t)i)); if (rc != PT_OK) return rc; }
    return PT_OK;
}
// n_frames on EVERY GPU (n_frames x size samples per pixel), one reduce to GPU 0, overlapped with the next call
int pt_group_render_batch(pt_group* g, int32_t n_frames) {
    if (!g || n_frames < 1) return fail(PT_ERR_INVALID, "pt_group_render_batch: bad argument");
    if (g->ctx.size() == 1) return pt_render_frames(g->ctx[0], n_frames);
    GROUP_EACH(g, batch_render(c, n_frames));
    NcclApi* api = nccl_api();
    NK(api->GroupStart());                         // one thread issues every rank's collective: they must be grouped
    int rc = PT_OK;
    for (pt_context* c : g->ctx) { int r = batch_reduce(c, n_frames); if (r != PT_OK) rc = r; }
    NK(api->GroupEnd());
    if (rc != PT_OK) return rc;
    GROUP_EACH(g, batch_finish(c, n_frames));
    return PT_OK;
}
int pt_group_reset_samples(pt_group* g) {
    if (!g) return fail(PT_ERR_INVALID, "pt_group_reset_samples: null group");
    if (g->ctx.size() == 1) return pt_reset_samples(g->ctx[0]);
    GROUP_EACH(g, pt_comm_reset(c));
    return PT_OK;
}
uint32_t pt_group_num_samples(pt_group* g) {
    if (!g || g->ctx.empty()) return 0;
    return g->ctx.size() == 1 ? pt_num_samples(g->ctx[0]) : pt_comm_num_samples(g->ctx[0]);
}
int pt_group_synchronize(pt_group* g) {
    if (!g) return fail(PT_ERR_INVALID, "pt_group_synchronize: null group");
    if (g->ctx.size() == 1) return pt_synchronize(g->ctx[0]);
    GROUP_EACH(g, pt_comm_synchronize(c));
    return PT_OK;
}
int pt_group_download_total(pt_group* g, float* rgba) {
    if (!g || !rgba) return fail(PT_ERR_INVALID, "pt_group_download_total: bad argument");
    return g->ctx.size() == 1 ? pt_download_accum(g->ctx[0], rgba) : pt_comm_download_total(g->ctx[0], rgba);
}
int pt_group_save_screenshot(pt_group* g, const char* path) {
    if (!g || !path) return fail(PT_ERR_INVALID, "pt_group_save_screenshot: bad argument");
    pt_context* c = g->ctx[0];
    if (g->ctx.size() == 1) return pt_save_screenshot(c, path);
    int rc = pt_group_synchronize(g);
    if (rc != PT_OK) return rc;
    // Renderer::SaveScreenshot of the combined image: the running total with its own sample count through the same writer
    float4* keep = c->d_accum; const uint32_t keep_n = c->num_samples;
    c->d_accum = c->comm->total.p; c->num_samples = c->comm->total_samples;
    rc = pt_save_screenshot(c, path);
    c->d_accum = keep; c->num_samples = keep_n;
    return rc;
}
int pt_group_present_async(pt_group* g, uint8_t* host_rgb_u8) {
    if (!g || !host_rgb_u8) return fail(PT_ERR_INVALID, "pt_group_present_async: bad argument");
    if (g->ctx.size() == 1) return pt_present_async(g->ctx[0], pt_num_samples(g->ctx[0]), host_rgb_u8);
    return pt_comm_present_async(g->ctx[0], host_rgb_u8);
}

}  // extern "C"
