// rt_internal.h — host-side state shared by the translation units of librtb200.so.
#ifndef RTB200_INTERNAL_H
#define RTB200_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "rtb200.h"
#include "rt_scene.cuh"

namespace rth {

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define RT_CUDA(call)                                                          \
    do {                                                                       \
        cudaError_t _e = (call);                                               \
        if (_e != cudaSuccess) return rth::cuda_fail(_e, #call, __FILE__, __LINE__); \
    } while (0)

// Grow-only device buffer (gputils::create_buffer / copy_to_gpu replacement, include/gputils/alloc.h:14-22:
// plain cudaMalloc, no managed memory; reused across frames instead of malloc/free per frame).
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);
    void release();
    template <class T> T* as() const { return (T*)p; }
};

struct Bvh {
    bool valid = false;
    int n_prims = 0;
    int n_bvh2_nodes = 0;
    int n_nodes8 = 0;
    int build_launches = 0;
    int build_path = 0;  // which builder produced the tree: 1 = multi-kernel pipeline, 2 = one-launch cluster kernel
    DevBuf prim_world;   // float4[3*n_prims]  world-space triangles in input (instance-major) order
    DevBuf prim_box;     // float[6*n_prims]   inflated prim AABBs
    DevBuf scene_bounds; // float[6] (+ atomics scratch)
    DevBuf keys_a, keys_b, vals_a, vals_b;  // radix sort ping-pong
    DevBuf hist;         // radix histograms
    DevBuf left, right, parent, range_lo, range_hi, flags;  // BVH2 topology
    DevBuf node_box;     // float[6*(2n-1)]
    DevBuf nodes8;       // uint4[5*n_nodes8]
    DevBuf prims8;       // float4[3*n_prims] in BVH8 leaf order
    DevBuf prim_order8;  // int[n_prims] sorted-order index per BVH8 prim slot
    const unsigned long long* sorted_keys = nullptr;  // into keys_a/keys_b after the last pass
    const int* sorted_vals = nullptr;
    DevBuf work;         // per-instance triangle prefix
    DevBuf collapse_scratch;  // level lists, slots, counts, block totals
    DevBuf rec;               // cluster build: 64-byte per-node records for the collapse
    DevBuf edge_box;          // cluster build: prefix / suffix unions of each CTA's leaf chunk
    int cluster_size = 0;     // CTAs of the last one-launch build
    // kept for rt_bvh_refit: which BVH2 node every BVH8 node / slot came from, and the instance layout the
    // topology was built for (refit is only valid while every instance keeps its mesh)
    DevBuf node8_root;   // int[n_nodes8]
    DevBuf node8_slot;   // int[8*n_nodes8]
    bool topology_valid = false;
    int n_levels8 = 0;   // depth of the BVH8 (finish_build)
    std::vector<int> topo_meshes;  // mesh index per instance at build time
    int* h_off = nullptr;          // pinned staging of the per-instance triangle prefix
    size_t h_off_cap = 0;
    bool off_in_flight = false;
    int prefix_uploads = 0;        // how many builds had to upload the prefix (tests: unchanged meshes upload once)
    void release();
};

}  // namespace rth

namespace rth {
extern unsigned long long g_alloc_gen;  // grows whenever a DevBuf is reallocated: device pointers baked into a captured frame are stale then

// What rt_frame_finish needs to know about an enqueued frame; saved with a captured frame and restored at every replay.
struct FrameState {
    bool pend_built = false, pend_wavefront = false, pend_stats = false, pend_direct = false;
    bool split_shade_used = false, tile_kernel_used = false, exchange_pending = false;
    int pend_waves = 0, pend_chunks = 1, wave0_form = 0, sb_used = -1, launches = 0;
};
// One instantiated CUDA graph of a frame (RT_RENDER_GRAPH) and the key it is valid for.
struct FrameGraph {
    std::vector<unsigned char> key;
    cudaGraphExec_t exec = nullptr;
    FrameState state;
    unsigned long long last_used = 0;
};
}  // namespace rth

struct RtScene {
    // host copies (instances can be replaced; the rest is immutable after create)
    std::vector<RtInstance> instances;
    std::vector<RtMesh> meshes;
    int n_vertices = 0, n_triangles = 0, n_meshes = 0, n_instances = 0, n_lights = 0, n_materials = 0;
    RtCamera camera;
    RtEnv env;
    int width = 0, height = 0;
    int any_refractive = 0, any_secondary = 0, cull_ok = 0;
    bool trace_unlit = false;       // RT_RENDER_TRACE_UNLIT
    std::vector<int> mesh_tri_count;
    // device
    rth::DevBuf d_tri_pre;          // float4 per triangle record: unit plane normal + area (k_tri_precompute)
    rth::DevBuf d_verts, d_norms, d_tris, d_meshes, d_insts, d_mats, d_lights, d_mesh_box;
    rth::DevBuf d_texc, d_atlas;    // TextureCoords per triangle record + RGBA atlas; empty when nothing is textured
    int atlas_w = 0, atlas_h = 0;
    bool textured = false;
    rth::DevBuf d_canvas, d_counters, d_hitq, d_paths, d_queue_a, d_queue_b, d_ids;
    rth::DevBuf d_shadowq, d_occl, d_shq_count;
    rth::DevBuf d_push_ctr;         // block counter of the signalled tile push
    // RT_RENDER_PEER_STORE (rt_scene_set_peers): [parity][rank] canvases, frame epoch, gathering rank (-1: every rank)
    std::vector<uint32_t*> peer_canv;
    int peer_n = 0, peer_root = -1;
    long long peer_pixels = 0;
    unsigned peer_epoch = 0;
    rth::DevBuf d_sort_keys_a, d_sort_keys_b, d_sort_vals_b, d_sort_hist;  // RT_RENDER_SORT_RAYS
    bool sort_rays = false;
    bool fused_shade = false;       // RT_RENDER_FUSED_SHADE: one k_shade kernel instead of gen / traverse / resolve
    bool split_shade_used = false;
    bool split_shade = false;       // RT_RENDER_SPLIT_SHADE: never use the one-kernel wave 0 (k_tile)
    bool tile_kernel_used = false;
    bool tile_kernel = false;       // RT_RENDER_TILE_KERNEL
    bool fast_accept = false;       // RT_RENDER_FAST_ACCEPT
    bool meshes_identity = true, all_identity = false;  // no rotated mesh / instance pose (tolerance mode needs it)
    // wave-0 shadow kernel auto-selection: frames timed so far, best k_shadow time per variant [per-lane, warp-coherent],
    // the variant the pending frame used (-1: not applicable), the choice (-1: still measuring)
    // One tuner per tile shape (tile size, world size, canvas size): frames with different tilings — e.g. the 32x16 tiles of a
    // device-resident frame and the 4-row full-width bands of a host frame — time the two kernels differently.
    struct SbTuner { int shape[5] = {0, 0, 0, 0, 0}; int frames = 0, choice = -1; float ms[3] = {0.0f, 0.0f, 0.0f}; bool used = false; };
    SbTuner tuners[6];
    int sb_cur = 0, sb_next = 0, sb_used = -1;
    bool pend_stats = false;
    int wave0_form = 0;             // 0 split, 1 tile kernels, 2 beam primary + split shading (RtFrameStats.tile_kernel)
    rth::Bvh bvh;
    bool boxes_valid = false;
    cudaStream_t stream = nullptr;      // the stream all work of this scene goes to: own_stream or the caller's (rt_scene_set_stream)
    cudaStream_t own_stream = nullptr;
    bool pending = false;               // a frame is enqueued and rt_frame_finish() has not run yet
    bool pend_built = false, pend_wavefront = false;
    int pend_waves = 0, pend_launches0 = 0;
    cudaEvent_t ev[12] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    // A frame rendered into a host canvas is cut into up to RT_MAX_CHUNKS chunks of tile rows; chunk c is copied to the
    // host on copy_stream while chunk c + 1 renders. Every chunk has its own counter block and events.
    static constexpr int kMaxChunks = 8;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t cev[kMaxChunks][6] = {};  // per chunk: start, primary done, shadow rays generated, traced, wave 0 shaded, waves done
    cudaEvent_t chunk_done[kMaxChunks] = {};  // no timing; copy_stream waits on these
    cudaEvent_t copies_done = nullptr;
    cudaEvent_t ev_exchange = nullptr;  // recorded behind the mailbox wait of a signalled exchange
    bool exchange_pending = false;      // ... whose outcome (h_pinned[48]) rt_frame_finish still has to look at
    int pend_chunks = 1;
    bool pend_direct = false;
    // Concurrent lanes (RT_FRAME_LANES, render_enqueue): a device-resident frame without secondary rays is cut into up to
    // kMaxChunks contiguous ranges of owned tiles; lane c > 0 runs its wave-0 kernels on lane_stream[c - 1], forked from and
    // joined back into the scene's stream, with its own counter block, events and slice of the hit / shadow queues, so that
    // the tail of one lane's kernel (a few SMs finishing their longest work units) is filled by the next lane's kernel.
    cudaStream_t lane_stream[kMaxChunks - 1] = {};
    cudaEvent_t lane_fork = nullptr, lane_join[kMaxChunks - 1] = {};
    cudaStream_t cur_stream = nullptr;  // stream of the lane run_wavefront is enqueuing (nullptr: the scene's stream)
    long long cur_off = 0;              // first hit-queue slot of that lane
    int cur_lane = 0, pend_lanes = 1;
    int lanes_req = 0;                  // rt_scene_set_lanes (0: RT_FRAME_LANES / 1)
    // pinned host memory, u64 units: [32,36) build result, [48] exchange time-out mask, [50,58) phase time stamps of the cluster build (RT_BUILD_TIMERS), [40,42) live counts of the bounce waves,
    // [64 + 32 c, 96 + 32 c) counters of chunk c
    unsigned long long* h_pinned = nullptr;
    void* h_inst = nullptr;             // pinned staging of the instance records (rt_scene_set_instances)
    size_t h_inst_cap = 0;
    int sm_count = 148;
    int kernel_launches = 0;  // running count, sampled around each API call
    // RT_RENDER_GRAPH: the frame is captured once per key (render flags, tile map, canvas, camera, environment, instance
    // generation, buffer generation, BVH state, exchange parity) and replayed with one cudaGraphLaunch
    bool capturing = false;              // the scene's stream is under capture: timing events become external record nodes
    rth::FrameGraph graphs[4];
    std::vector<unsigned char> last_key, prev_key; // keys of the last two frames that asked for a graph (a frame is captured on its second appearance)
    std::vector<unsigned char> bad_key;  // a key whose capture failed: plain launches from then on
    unsigned long long graph_clock = 0, inst_gen = 0;
    int graph_replays = 0, graph_captures = 0, pend_graph = 0;
};

namespace rth {
// sync = false: everything is enqueued, finish_build() must run after the next stream synchronisation
int build_bvh(RtScene* sc, RtBuildStats* stats, bool sync = true);   // rt_bvh_build.cu
int finish_build(RtScene* sc, RtBuildStats* stats);
// keeps the tree, recomputes triangles and every box (instances moved, none added / removed / re-meshed)
int refit_bvh(RtScene* sc, RtBuildStats* stats, bool sync = true);
bool refit_possible(const RtScene* sc);
// stable LSD radix sort of (u64 key, int value) pairs on the low 8 * n_passes key bits; the a pointers end up at the result
size_t radix_hist_bytes(int n);
int radix_sort_pairs(cudaStream_t st, unsigned long long*& keys_a, unsigned long long*& keys_b, int*& vals_a, int*& vals_b,
                     unsigned* hist, int n, int n_passes);
int update_instance_boxes(RtScene* sc);                    // rt_engine.cu
// cudaEventRecord that stays usable for timing when the stream is being captured into a graph
inline cudaError_t record(const RtScene* sc, cudaEvent_t ev, cudaStream_t st) {
    return cudaEventRecordWithFlags(ev, st, sc->capturing ? cudaEventRecordExternal : cudaEventRecordDefault);
}
rtd::SceneView make_view(const RtScene* sc, bool with_bvh);
}  // namespace rth

#endif
