// Batch: packs worlds into HBM blobs, launches the step kernels, moves state in and out.
// Compiled with:  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo --fmad=false
#include "engine.h"

#include <algorithm>
#include <atomic>
#include <exception>
#include <thread>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>

#include "kernels.cuh"
#include "wkernel.cuh"
#include "gkernel.cuh"

namespace mevo {

#define CUDA_CHECK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                                     __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");   \
    } while (0)

namespace {

// size classes: 0, 1, 2 = warp-per-environment kernel (wkernel.cuh) with compile-time capacities;
// 3..6 = CTA-per-environment kernel (gkernel.cuh) with 64 / 128 / 256 / 512 threads per environment
constexpr int kNumClasses = 7;
constexpr int kFirstGroupClass = 3;
constexpr int kLastClass = kNumClasses - 1;
const int kGroupSizes[kNumClasses] = {32, 32, 32, 64, 128, 256, 512};
constexpr int kMaxSmemPerBlock = 227 * 1024;
constexpr int kGroupMaxPoints = 1024, kGroupMaxTypes = 64;      // 10-bit point ids, 6-bit edge types in the packed edge word
// warp-kernel capacities: points, edges, leaves, tested pairs, always-items, rigid-cell visits
#define MEVO_WCAPS_A 40, 112, 24, 192, 256, 112     /* one 5x5 robot on static terrain (Walker-v0, Climber-v0, ...) */
#define MEVO_WCAPS_C 52, 144, 28, 320, 288, 64      /* 5x5 robot + the 3x2 package (Pusher-v0, Thrower-v0): 18.6 KB, 12 per SM */
#define MEVO_WCAPS_B 64, 160, 32, 512, 640, 128     /* 5x5 robot + a larger free object (Carrier, Catcher, beams): 22.5 KB, 10 per SM */

int group_class_for(int Pd, int Ed) {
    // threads per environment.  MEVO_GCLASS = "p64,p128,p256" (development): largest point count that runs with 64 / 128 / 256
    // threads; MEVO_GROUP_SHIFT (development, tests) moves every environment up or down one class.  A class whose largest
    // environment has more points than threads runs a 2- or 4-points-per-thread build of the kernel.
    int lim[3] = {128, 192, 512};          // measured on B200: few threads per environment, many environments per SM (the middle
                                           // limit from a sweep on the heterogeneous mix: 256 -> 393 k, 160 ... 208 -> 399 ... 405 k env-steps/s)
    if (const char* g = std::getenv("MEVO_GCLASS")) std::sscanf(g, "%d,%d,%d", &lim[0], &lim[1], &lim[2]);
    const char* shift_env = std::getenv("MEVO_GROUP_SHIFT");       // read at every finalize: the tests move fixtures across classes
    const int shift = shift_env ? std::atoi(shift_env) : 0;
    (void)Ed;
    int c = kLastClass;
    if (Pd <= lim[0]) c = kFirstGroupClass;
    else if (Pd <= lim[1]) c = kFirstGroupClass + 1;
    else if (Pd <= lim[2]) c = kFirstGroupClass + 2;
    c = shift ? std::min(kLastClass, std::max(kFirstGroupClass, c + shift)) : c;
    while (c < kLastClass && Pd > 4 * kGroupSizes[c]) c++;          // at most four points per thread
    return c;
}

int class_for(const DesignHeader& H, bool use_warp_kernel) {
    if (use_warp_kernel && H.warp_ok && H.max_deg <= 8 && H.unit_mass && H.n_obj <= 8 && H.n_types <= 16) {
        const int caps[3][6] = {{MEVO_WCAPS_A}, {MEVO_WCAPS_C}, {MEVO_WCAPS_B}};
        for (int c = 0; c < 3; c++)
            if (H.Pd <= caps[c][0] && H.Ed <= caps[c][1] && H.n_leaf <= caps[c][2] && H.n_pairs <= caps[c][3] &&
                H.n_items <= caps[c][4] && H.n_rvis <= caps[c][5]) return c;
    }
    return group_class_for(H.Pd, H.Ed);
}

uint64_t fnv1a(const std::vector<uint8_t>& v) {
    uint64_t h = 1469598103934665603ull;
    for (uint8_t b : v) { h ^= b; h *= 1099511628211ull; }
    return h;
}

struct BlobWriter {
    std::vector<uint8_t>& buf;
    explicit BlobWriter(std::vector<uint8_t>& b) : buf(b) {}
    int begin_section() {
        while (buf.size() % 16) buf.push_back(0);
        return (int)buf.size();
    }
    template <class T>
    void put(const T& v) {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(&v);
        buf.insert(buf.end(), p, p + sizeof(T));
    }
    template <class T>
    void put_array(const std::vector<T>& v) {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
        buf.insert(buf.end(), p, p + sizeof(T) * v.size());
    }
};

// environments per CTA that make the launch a whole number of equally full waves (slots = CTAs resident at once)
int balanced_envs_per_cta(int n_envs, int epc, int slots) {
    const int ctas = (n_envs + epc - 1) / epc;
    const int waves = (ctas + slots - 1) / slots;
    const int per = (n_envs + waves * slots - 1) / (waves * slots);
    return std::max(1, std::min(epc, per));
}

// Kernel attributes are per device AND shared by every batch of the process on that device: every Batch::finalize raises the
// dynamic shared-memory limit of the kernels its classes use, on its own device, to what it needs -- never lowers it, or a batch
// finalized later with smaller robots would make an earlier batch's launches fail ("invalid argument").
template <class Kernel>
void raise_dynamic_smem(Kernel kernel, int bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, int> limit;
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    int& cur = limit[{dev, reinterpret_cast<const void*>(kernel)}];
    if (bytes > cur) {
        CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        cur = bytes;
    }
}

template <int PCAP, int ECAP, int LCAP, int PAIRCAP, int ITEMCAP, int VISCAP, int EPC, int MINB>
void configure_wstep() {
    constexpr int smem = WarpSmem<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP>::kBytes * EPC;
    static_assert(smem <= kMaxSmemPerBlock, "warp kernel CTA does not fit in shared memory");
    raise_dynamic_smem(wstep_kernel<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, false>, smem);
    raise_dynamic_smem(wstep_kernel<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, true>, smem);
}

template <int PCAP, int ECAP, int LCAP, int PAIRCAP, int ITEMCAP, int VISCAP, int EPC, int MINB>
void launch_wstep(StepParams prm, int n_sm, bool balance, cudaStream_t stream) {
    using Smem = WarpSmem<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP>;
    if (prm.work) {
        // one CTA of the full width per (CTA-group, chunk) item, drawn from a counter as the CTA starts (wkernel.cuh); the
        // counters start at zero
        prm.envs_per_cta = EPC;
        const int groups = (prm.n_class_envs + EPC - 1) / EPC;
        CUDA_CHECK(cudaMemsetAsync(prm.work, 0, sizeof(int) * (size_t)(1 + groups), stream));
        const int grid = groups * prm.n_chunks;
        wstep_kernel<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, true><<<grid, 32 * EPC, Smem::kBytes * EPC, stream>>>(prm);
        return;
    }
    prm.envs_per_cta = balance ? balanced_envs_per_cta(prm.n_class_envs, EPC, n_sm * MINB) : EPC;
    // shared memory for the environments the CTA really takes: what is left of the SM's 228 KB serves as L1 for the static-terrain
    // tables (grid, boxes, leaf records, points), which the collision phase reads through __ldg
    const int smem = Smem::kBytes * prm.envs_per_cta;
    const int grid = (prm.n_class_envs + prm.envs_per_cta - 1) / prm.envs_per_cta;
    wstep_kernel<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, false><<<grid, 32 * EPC, smem, stream>>>(prm);
}

// group kernel builds: (threads, CTAs per SM the register budget is set for, points per thread)
enum class GAction { Configure, Launch };
template <int G, int MINB, int PPT, bool XD>
void gstep_do2(GAction what, const StepParams* prm, int smem, cudaStream_t stream) {
    if (what == GAction::Configure)
        raise_dynamic_smem(gstep_kernel<G, MINB, PPT, XD>, smem);
    else
        gstep_kernel<G, MINB, PPT, XD><<<prm->n_class_envs, G, smem, stream>>>(*prm);
}
template <int G, int MINB, int PPT>
void gstep_do(bool xd, GAction what, const StepParams* prm, int smem, cudaStream_t stream) {
    if (xd) gstep_do2<G, MINB, PPT, true>(what, prm, smem, stream);
    else gstep_do2<G, MINB, PPT, false>(what, prm, smem, stream);
}

constexpr int kPptExtBit = 8;                 // class_ppt flag: a design of the class has gather slots 8..15 (DesignHeader::o_csrx)

void gstep_dispatch(int cls, int ppt_x, GAction what, const StepParams* prm, int smem, cudaStream_t stream) {
    const int g = kGroupSizes[cls];
    const bool xd = (ppt_x & kPptExtBit) != 0;
    const int ppt = ppt_x & ~kPptExtBit;
    if (ppt > 2) {
        if (g == 64) gstep_do<64, 8, 4>(xd, what, prm, smem, stream);
        else if (g == 128) gstep_do<128, 4, 4>(xd, what, prm, smem, stream);
        else if (g == 256) gstep_do<256, 2, 4>(xd, what, prm, smem, stream);
        else gstep_do<512, 1, 4>(xd, what, prm, smem, stream);
    } else if (ppt == 2) {
        if (g == 64) gstep_do<64, 8, 2>(xd, what, prm, smem, stream);
        else if (g == 128) gstep_do<128, 4, 2>(xd, what, prm, smem, stream);
        else if (g == 256) gstep_do<256, 2, 2>(xd, what, prm, smem, stream);
        else gstep_do<512, 1, 2>(xd, what, prm, smem, stream);
    } else {
        if (g == 64) gstep_do<64, 8, 1>(xd, what, prm, smem, stream);
        else if (g == 128) gstep_do<128, 4, 1>(xd, what, prm, smem, stream);
        else if (g == 256) gstep_do<256, 2, 1>(xd, what, prm, smem, stream);
        else gstep_do<512, 1, 1>(xd, what, prm, smem, stream);
    }
}

// ---- allocation-free stepping (Batch::step_static_scratch).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The group kernel's neighbour-list scratch lives in this file-scope device array instead of the
// cudaMalloc'ed block Batch::finalize sizes: kStaticScratchBytes (32 MiB, ~1000 environments of 20x20 robots) are held by
// every GPU that loads the library until the process exits.  In HOST code the array's name is only a stand-in for the
// symbol, not a device address; the address a kernel can use comes from cudaGetSymbolAddress, resolved once per device
// with that device current.  All calls share the array and it keeps the last call's contents; nothing in it has to start
// at zero (the kernel rebuilds every list it reads in the first substep of each launch, their lengths live in shared
// memory), but two calls that overlap on one GPU would race on it.
constexpr size_t kStaticScratchBytes = 32u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;
std::mutex g_static_scratch_mutex;
unsigned int* g_static_scratch_ptr[kMaxDevices] = {};

// device address of g_static_scratch on `device` (which must be current)
unsigned int* static_scratch_address(int device) {
    if (device < 0 || device >= kMaxDevices) throw std::runtime_error("static scratch: device index out of range");
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_ptr[device]) {
        void* ptr = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) throw std::runtime_error("static scratch: cudaGetSymbolAddress returned a null address");
        g_static_scratch_ptr[device] = static_cast<unsigned int*>(ptr);
    }
    return g_static_scratch_ptr[device];
}

__global__ void reset_kernel(double* state, const double* state0, const EnvDesc* envs, const uint8_t* design,
                             const int* env_ids, int n, int* sim_time, uint8_t* unstable) {
    // Environment::revert_to_snapshot(0) (Environment.cpp:247-263): pos, vel <- snapshot; pos_last <- pos;
    // vel_true <- vel; length_eq_goal is NOT touched.
    const int k = blockIdx.x;
    if (k >= n) return;
    const int env = env_ids ? env_ids[k] : k;
    const EnvDesc ed = envs[env];
    const int Pd = reinterpret_cast<const DesignHeader*>(design + ed.design_off)->Pd;
    double2* s = reinterpret_cast<double2*>(state + ed.state_off);
    const double2* s0 = reinterpret_cast<const double2*>(state0 + ed.state_off);
    for (int i = threadIdx.x; i < Pd; i += blockDim.x) {
        const double2 p = s0[i], v = s0[Pd + i];
        s[i] = p; s[Pd + i] = v; s[2 * Pd + i] = p; s[3 * Pd + i] = v;
    }
    if (threadIdx.x == 0) { sim_time[env] = 0; unstable[env] = 0; }
}

// Environment::revert_to_snapshot(t) (Environment.cpp:247-263) from a device-resident snapshot slot: pos, vel <- slot;
// pos_last <- pos; vel_true <- vel; sim_time <- the slot's time; length_eq_goal is NOT touched.
__global__ void revert_kernel(double* state, const double* slot, const int* slot_time, const EnvDesc* envs, const uint8_t* design,
                              const int* env_ids, int n, int* sim_time, uint8_t* unstable) {
    const int k = blockIdx.x;
    if (k >= n) return;
    const int env = env_ids ? env_ids[k] : k;
    const EnvDesc ed = envs[env];
    const int Pd = reinterpret_cast<const DesignHeader*>(design + ed.design_off)->Pd;
    double2* s = reinterpret_cast<double2*>(state + ed.state_off);
    const double2* s0 = reinterpret_cast<const double2*>(slot + ed.state_off);
    for (int i = threadIdx.x; i < Pd; i += blockDim.x) {
        const double2 p = s0[i], v = s0[Pd + i];
        s[i] = p; s[Pd + i] = v; s[2 * Pd + i] = p; s[3 * Pd + i] = v;
    }
    if (threadIdx.x == 0) { sim_time[env] = slot_time[env]; unstable[env] = 0; }
}

}  // namespace

struct Batch::Impl {
    std::vector<std::vector<int>> class_envs;     // per class: env ids
    std::vector<int> class_smem;                  // per class: bytes of shared memory per env
    std::vector<size_t> class_off;                // per class: offset into d_class_envs_
    std::vector<EnvDesc> desc;
    std::vector<std::vector<double>> pos0, vel0;  // reset snapshot (global numbering)
    std::vector<int> env_cost;                    // per env: estimated work of one substep
    bool use_warp_kernel = true;
    int wepc = 0;                                 // warp kernel: environments per CTA override (0 = default)
    int class_ppt[kNumClasses] = {};              // group kernel: points per thread of the class's build (1, 2 or 4)
    int class_xd[kNumClasses] = {};               // kPptExtBit if a design of the class carries gather slots 8..15, else 0
    unsigned int* d_scratch = nullptr;            // group kernel: neighbour lists
    size_t scratch_bytes = 0;                     // bytes of neighbour-list scratch this batch's environments address
    // device-resident snapshot slots (Environment::save_snapshot / revert_to_snapshot, batched): a copy of the state buffer and of
    // the per-environment times each, allocated on first use of the slot
    static constexpr int kSlots = 8;
    double* d_slot[kSlots] = {};
    int* d_slot_time[kSlots] = {};
    double* d_cf = nullptr;                       // test hook: collision_forces of the first substep
    int cta_sync = 1;
    int balance = 1;                              // warp kernel: environments per CTA chosen for equally full waves
    int n_sm = 148;
    // heterogeneous batches: the size classes of one step run concurrently, each on its own stream, forked from and joined
    // to the batch's stream with events (largest environments first: they set the length of the step)
    cudaStream_t class_stream[kNumClasses] = {};
    cudaEvent_t class_done[kNumClasses] = {};
    cudaEvent_t fork = nullptr;
    int concurrent_classes = 1;
    // group classes: an env-step issued as `chunks` launches of consecutive substep ranges on `lanes` streams per class, each lane
    // a contiguous share of the class.  A lane's next chunk starts as soon as its own previous chunk is done, so the SMs that the
    // last, partly filled wave of one launch leaves idle run the other lanes' CTAs (1024 20x20 robots on 296 CTA slots are
    // 3.46 waves: 13 % of the slots idle in a single launch).  MEVO_CHUNKS / MEVO_LANES; 1 / 1 = one launch per class.
    int chunks = 4, lanes = 2;                    // most chunks considered; streams per class
    double chunk_min_waves = 1.0;                 // chunk only when the class asks for more CTAs than fit at once (MEVO_CHUNK_MIN_WAVES) ...
    double chunk_min_gain = 0.02;                 // ... and the estimated gain is at least this (MEVO_CHUNK_MIN_GAIN)
    static constexpr int kMaxLanes = 4;
    int wchunks = 6;                              // warp classes: most chunks considered for the item form (MEVO_WCHUNKS)
    int* d_work = nullptr;                        // its item counters, per warp class at work_off[c]
    int work_off[kNumClasses] = {};
    cudaStream_t lane_stream[kNumClasses][kMaxLanes] = {};
    cudaEvent_t lane_done[kNumClasses][kMaxLanes] = {};
    int class_slots[kNumClasses] = {};            // CTAs of the class's build that fit on the device at once
    int class_max[kNumClasses][6] = {};
    int* d_ids = nullptr;                         // reset_envs: environment ids on the device
    int ids_cap = 0;
    int* d_com_ranges = nullptr;
    double* d_com_out = nullptr;
    int com_obj = -1;
    uint8_t* unstable_host = nullptr;             // pinned staging buffer of the per-step flags (n bytes)
    std::vector<int> time_host;
    bool time_dirty = true;
};

Batch::Batch(int device) : device_(device), impl_(new Impl()) {
    CUDA_CHECK(cudaSetDevice(device_));
    cudaStream_t s;
    CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    stream_ = s;
    if (const char* e = std::getenv("MEVO_NO_WKERNEL")) impl_->use_warp_kernel = std::atoi(e) == 0;
    if (const char* e = std::getenv("MEVO_WEPC")) impl_->wepc = std::atoi(e);
    if (const char* e = std::getenv("MEVO_CTA_SYNC")) impl_->cta_sync = std::atoi(e);
    if (const char* e = std::getenv("MEVO_BALANCE")) impl_->balance = std::atoi(e);
    if (const char* e = std::getenv("MEVO_CONCURRENT_CLASSES")) impl_->concurrent_classes = std::atoi(e);
    if (const char* e = std::getenv("MEVO_CHUNKS")) impl_->chunks = std::max(1, std::atoi(e));
    if (const char* e = std::getenv("MEVO_CHUNK_MIN_WAVES")) impl_->chunk_min_waves = std::atof(e);
    if (const char* e = std::getenv("MEVO_CHUNK_MIN_GAIN")) impl_->chunk_min_gain = std::atof(e);
    if (const char* e = std::getenv("MEVO_WCHUNKS")) impl_->wchunks = std::max(1, std::atoi(e));
    if (const char* e = std::getenv("MEVO_LANES")) impl_->lanes = std::min((int)Impl::kMaxLanes, std::max(1, std::atoi(e)));
    CUDA_CHECK(cudaDeviceGetAttribute(&impl_->n_sm, cudaDevAttrMultiProcessorCount, device_));
}

Batch::~Batch() {
    cudaSetDevice(device_);
    task_free();
    sight_free();
    for (void* g : graphs_) if (g) cudaGraphExecDestroy((cudaGraphExec_t)g);
#ifdef MEVO_PROF
    {
        unsigned long long gp[8];
        if (cudaMemcpyFromSymbol(gp, g_g2prof, sizeof(gp)) == cudaSuccess) {
            unsigned long long tot = 0;
            for (int q = 0; q < 8; q++) tot += gp[q];
            if (tot) {
                const char* names[8] = {"boxes+list", "broad+filter", "resolve+contacts", "rk4", "band test", "jacobi", "rigid pass", "true vel"};
                std::fprintf(stderr, "[prof] group kernel, share of CTA cycles:");
                for (int q = 0; q < 8; q++) std::fprintf(stderr, " %s %.3f", names[q], (double)gp[q] / tot);
                std::fprintf(stderr, "\n");
                unsigned long long gc[8];
                if (cudaMemcpyFromSymbol(gc, g_g2cnt, sizeof(gc)) == cudaSuccess && gc[0]) {
                    const double n = (double)gc[0];
                    std::fprintf(stderr, "[prof] per CTA-substep: list rebuilds %.3f survivors %.2f contact batches %.3f contact rounds %.3f jacobi %.3f "
                                 "rigid pass %.3f rigid moves %.2f\n", gc[1] / n, gc[2] / n, gc[3] / n, gc[4] / n, gc[5] / n, gc[6] / n, gc[7] / n);
                }
            }
        }
    }
    {
        unsigned long long h[16];
        if (cudaMemcpyFromSymbol(h, g_wprof, sizeof(h)) == cudaSuccess && h[3]) {
            const char* names[4] = {"collide", "rk4", "constrain", "substep"};
            for (int q = 0; q < 4; q++)
                std::fprintf(stderr, "[prof] %-9s own cycles share %.3f, slowest-warp / mean = %.3f\n", names[q], (double)h[q] / h[3], (double)h[4 + q] / h[q]);
            unsigned long long hist[3][16][2];
            cudaMemcpyFromSymbol(hist, g_whist, sizeof(hist));
            const char* hn[3] = {"survivors/4", "candidates/4", "constrain: rigid moves (+8: Jacobi ran)"};
            for (int t = 0; t < 3; t++) {
                std::fprintf(stderr, "[prof] %s cycles by %s:", t < 2 ? "collide" : "limiter", hn[t]);
                for (int b = 0; b < 16; b++)
                    if (hist[t][b][1]) std::fprintf(stderr, " %d:%.0f(n=%llu)", b, (double)hist[t][b][0] / hist[t][b][1], hist[t][b][1]);
                std::fprintf(stderr, "\n");
            }
        }
    }
#endif
    cudaFree(d_design_); cudaFree(d_static_); cudaFree(d_state_); cudaFree(d_state0_); cudaFree(d_act_);
    cudaFree(d_envdesc_); cudaFree(d_class_envs_); cudaFree(d_unstable_); cudaFree(d_time_); cudaFree(d_err_);
    cudaFree(d_cand_); cudaFree(d_cand_n_); cudaFree(d_events_);
    if (impl_) {
        cudaFree(impl_->d_com_ranges); cudaFree(impl_->d_com_out); cudaFree(impl_->d_ids);
        cudaFree(impl_->d_scratch); cudaFree(impl_->d_cf); cudaFree(impl_->d_work);
        if (impl_->unstable_host) cudaFreeHost(impl_->unstable_host);
        for (int k = 0; k < Impl::kSlots; k++) { cudaFree(impl_->d_slot[k]); cudaFree(impl_->d_slot_time[k]); }
        for (int c = 0; c < kNumClasses; c++) {
            if (impl_->class_stream[c]) cudaStreamDestroy(impl_->class_stream[c]);
            if (impl_->class_done[c]) cudaEventDestroy(impl_->class_done[c]);
            for (int l = 0; l < Impl::kMaxLanes; l++) {
                if (impl_->lane_stream[c][l]) cudaStreamDestroy(impl_->lane_stream[c][l]);
                if (impl_->lane_done[c][l]) cudaEventDestroy(impl_->lane_done[c][l]);
            }
        }
        if (impl_->fork) cudaEventDestroy(impl_->fork);
    }
    if (stream_) cudaStreamDestroy((cudaStream_t)stream_);
    delete impl_;
}

int Batch::begin_env(int world_w, int world_h) {
    envs_.emplace_back();
    envs_.back().world.init(world_w, world_h);
    finalized_ = false;
    return (int)envs_.size() - 1;
}

bool Batch::add_object(int env, const double* grid, int w, int h, const double* conn, int k,
                       const std::string& name, double x, double y, bool is_robot) {
    EnvHost& e = envs_.at(env);
    if (finalized_) download_all();
    const int P_before = e.world.num_points(), E_before = e.world.num_edges();
    const size_t objects_before = e.world.objects.size();
    if (!e.world.add_object(grid, w, h, conn, k, name, x, y, is_robot)) return false;
    if (e.world.objects.size() == objects_before) return true;       // a grid without voxels: success, nothing added (like the reference)
    // extend the host mirror with the new object's rest state (Environment.cpp:44-109)
    const World& W = e.world;
    e.pos.insert(e.pos.end(), W.pos.begin() + 2 * P_before, W.pos.end());
    e.pos_last.insert(e.pos_last.end(), W.pos.begin() + 2 * P_before, W.pos.end());
    e.vel.resize(2 * W.num_points(), 0.0);
    e.vel_true.resize(2 * W.num_points(), 0.0);
    e.goal.insert(e.goal.end(), W.el0.begin() + E_before, W.el0.end());
    e.obj_force_dynamic.push_back(0);
    e.host_fresh = true;
    if (finalized_) finalized_ = false;       // (read-only when already false: mevo_batch_add_robot_envs fills environments from several threads)
    return true;
}

// ------------------------------------------------------------------------------------------------
// packing

void Batch::pack_env(int env_id, std::vector<uint8_t>& design, std::vector<uint8_t>& statics) {
    EnvHost& e = envs_[env_id];
    const World& W = e.world;
    const int P = W.num_points(), E = W.num_edges(), n_obj = (int)W.objects.size();

    std::vector<uint8_t> dyn(n_obj, 0);
    for (int o = 0; o < n_obj; o++)
        dyn[o] = !(W.objects[o].all_fixed && !W.objects[o].is_robot) || e.obj_force_dynamic[o];

    e.point_local.assign(P, 0);
    e.edge_local.assign(E, -1);
    e.dyn_points.clear(); e.dyn_edges.clear();
    std::vector<int> static_points;
    for (int o = 0; o < n_obj; o++) {
        const VoxelObject& ob = W.objects[o];
        for (int p = ob.min_point; p <= ob.max_point; p++) {
            if (dyn[o]) { e.point_local[p] = (int)e.dyn_points.size(); e.dyn_points.push_back(p); }
            else { e.point_local[p] = -(int)static_points.size() - 1; static_points.push_back(p); }
        }
        if (dyn[o])
            for (int k = ob.min_edge; k <= ob.max_edge; k++) { e.edge_local[k] = (int)e.dyn_edges.size(); e.dyn_edges.push_back(k); }
    }
    const int Pd = (int)e.dyn_points.size(), Ed = (int)e.dyn_edges.size();
    if (Pd > kGroupMaxPoints || Ed > 32000)
        throw std::runtime_error("environment too large: more than 1024 simulated points (or 32000 springs) in one environment");

    auto make_leaf = [&](const VoxelObject& ob, const Boxel& b, bool is_dyn, int obj_slot) {
        LeafRec r;
        std::memset(&r, 0, sizeof(r));
        for (int q = 0; q < 4; q++) {
            const int lp = e.point_local[b.pt[q]];
            r.pt[q] = (uint16_t)(is_dyn ? lp : -lp - 1);
            const int ge = b.ed[q];
            const int la = e.point_local[W.ea[ge]], lb = e.point_local[W.eb[ge]];
            r.ea[q] = (uint16_t)(is_dyn ? la : -la - 1);
            r.eb[q] = (uint16_t)(is_dyn ? lb : -lb - 1);
            if (W.on_surface[ge]) r.surf_mask |= (uint8_t)(1u << q);
        }
        r.obj = (uint8_t)obj_slot;
        r.gx = (int16_t)(b.grid_index % ob.grid_w);
        r.gy = (int16_t)(b.grid_index / ob.grid_w);
        return r;
    };

    // ---- dynamic design blob
    design.clear();
    BlobWriter dw(design);
    DesignHeader H;
    std::memset(&H, 0, sizeof(H));
    design.resize(sizeof(DesignHeader));
    H.Pd = Pd; H.Ed = Ed;

    std::vector<double> mass(Pd), ek(Ed), el0(Ed);
    std::vector<uint16_t> ea(Ed), eb(Ed);
    std::vector<uint8_t> fixed(Pd);
    for (int i = 0; i < Pd; i++) { mass[i] = W.mass[e.dyn_points[i]]; fixed[i] = W.fixed[e.dyn_points[i]]; }
    for (int k = 0; k < Ed; k++) {
        const int g = e.dyn_edges[k];
        ek[k] = W.ek[g]; el0[k] = W.el0[g];
        ea[k] = (uint16_t)e.point_local[W.ea[g]];
        eb[k] = (uint16_t)e.point_local[W.eb[g]];
    }
    // action slots: robots in load order, actuators in Robot::init order
    std::vector<int16_t> eact(2 * (size_t)Ed, -1);
    e.act_slot_obj.clear(); e.act_slot_boxel.clear();
    for (int o = 0; o < n_obj; o++) {
        const VoxelObject& ob = W.objects[o];
        if (!ob.is_robot) continue;
        for (int gidx : ob.actuator_grid_index) {
            const int bi = ob.actuator_grid_to_boxel.at(gidx);
            const Boxel& b = ob.boxels[bi];
            const int slot = (int)e.act_slot_obj.size();
            e.act_slot_obj.push_back(o); e.act_slot_boxel.push_back(bi);
            // H actuator drives top + bot, V actuator left + right (PhysicsEngine.cpp:47-68); ed = top,right,bot,left
            const int e0 = b.cell_type == CELL_ACT_H ? b.ed[0] : b.ed[3];
            const int e1 = b.cell_type == CELL_ACT_H ? b.ed[2] : b.ed[1];
            for (int ge : {e0, e1}) {
                const int le = e.edge_local[ge];
                if (le < 0) continue;
                if (eact[2 * le] < 0) eact[2 * le] = (int16_t)slot;
                else if (eact[2 * le + 1] < 0) eact[2 * le + 1] = (int16_t)slot;
            }
        }
    }
    e.n_act = (int)e.act_slot_obj.size();
    H.n_act = e.n_act;
    // per-point gather lists: a-sides by edge id, then b-sides by edge id (PhysicsEngine.cpp:530-531)
    std::vector<uint16_t> csro(Pd + 1, 0), csre;
    {
        std::vector<std::vector<uint16_t>> lists(Pd);
        for (int k = 0; k < Ed; k++) lists[ea[k]].push_back((uint16_t)(k << 1));
        for (int k = 0; k < Ed; k++) lists[eb[k]].push_back((uint16_t)((k << 1) | 1));
        for (int i = 0; i < Pd; i++) {
            csro[i] = (uint16_t)csre.size();
            csre.insert(csre.end(), lists[i].begin(), lists[i].end());
        }
        csro[Pd] = (uint16_t)csre.size();
    }
    // warp kernel: one 16-byte record per point (8 x u16 entries (edge << 1 | side), same order, 0xffff = empty)
    // and one u32 per edge (a | b << 8 | flags << 16)
    std::vector<uint16_t> csr8(8 * (size_t)Pd, 0xffff);
    int max_deg = 0;
    for (int i = 0; i < Pd; i++) {
        const int deg = csro[i + 1] - csro[i];
        max_deg = std::max(max_deg, deg);
        for (int j = 0; j < deg && j < 8; j++) csr8[8 * (size_t)i + j] = csre[csro[i] + j];
    }
    H.max_deg = max_deg;
    H.unit_mass = 1;
    for (int i = 0; i < Pd; i++) if (mass[i] != 1.0) H.unit_mass = 0;
    // both kernels rely on what the reference's world builder guarantees (ObjectCreator.h:36-47: mass 1.0).  A lattice point
    // touches at most 4 sides + 4 diagonals when neighbouring voxels are connected; between UNCONNECTED neighbours the side
    // edge exists once per voxel, up to 4 more: those designs carry a second record per point (slots 8..15, group kernel only)
    if (!H.unit_mass) throw std::runtime_error("point masses other than 1.0 are not supported");
    if (max_deg > 16) throw std::runtime_error("a point with more than 16 incident springs is not supported");
    std::vector<uint16_t> csrx;
    if (max_deg > 8) {
        csrx.assign(8 * (size_t)Pd, 0xffff);
        for (int i = 0; i < Pd; i++)
            for (int j = 8; j < csro[i + 1] - csro[i]; j++) csrx[8 * (size_t)i + j - 8] = csre[csro[i] + j];
    }
    // leaves, objects, pair table
    std::vector<LeafRec> leaves;
    std::vector<ObjRec> objrecs;
    std::vector<int> leaf_obj_slot;
    for (int o = 0; o < n_obj; o++) {
        if (!dyn[o]) continue;
        const VoxelObject& ob = W.objects[o];
        ObjRec r;
        std::memset(&r, 0, sizeof(r));
        r.is_robot = ob.is_robot ? 1 : 0;
        r.n_surface = (uint16_t)ob.surface_boxels.size();
        r.leaf_start = (uint16_t)leaves.size();
        const int slot = (int)objrecs.size();
        for (int bi : ob.surface_boxels) { leaves.push_back(make_leaf(ob, ob.boxels[bi], true, slot)); leaf_obj_slot.push_back(slot); }
        r.leaf_count = (uint16_t)(leaves.size() - r.leaf_start);
        objrecs.push_back(r);
    }
    // Per-object device state is the self-collision counter of ROBOTS only (PhysicsEngine.cpp:856-863): robots, which load first
    // (evogym/sim.py:52-62), must sit in the first kMaxDynObjects slots; other simulated objects (the 27 blocks of Traverser-v0)
    // only need their slot to fit the pair code's object field.  More than kMaxDynObjects objects run on the group kernels.
    if ((int)objrecs.size() > kMaxPairObjects) throw std::runtime_error("too many dynamic objects in one environment (max 64)");
    for (size_t i = kMaxDynObjects; i < objrecs.size(); i++)
        if (objrecs[i].is_robot) throw std::runtime_error("too many robots in one environment (max 8, loaded before the other objects)");
    if ((int)leaves.size() > kMaxLeaves) throw std::runtime_error("too many surface voxels in one environment (max 4095)");
    // Leaf pairs inside one object that share a point ALWAYS have overlapping boxes (both boxes contain
    // the shared point), so they are candidates in every substep: they are not tested; their non-shared
    // corners are listed once as narrow-phase items.  Every other pair goes to the tested pair table.
    std::vector<uint32_t> pairs, items;
    e.always_pairs.clear();
    auto shares_point = [&](const LeafRec& x, const LeafRec& y) {
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++)
                if (x.pt[i] == y.pt[j]) return true;
        return false;
    };
    for (size_t a = 0; a < leaves.size(); a++)
        for (size_t b = a + 1; b < leaves.size(); b++) {
            const bool same = leaf_obj_slot[a] == leaf_obj_slot[b];
            if (same && shares_point(leaves[a], leaves[b])) {
                e.always_pairs.push_back((int)a); e.always_pairs.push_back((int)b);
                for (int dir = 0; dir < 2; dir++) {
                    const size_t m = dir ? b : a, r = dir ? a : b;
                    for (int c = 0; c < 4; c++) {
                        bool shared = false;
                        for (int j = 0; j < 4; j++) shared |= leaves[m].pt[c] == leaves[r].pt[j];
                        if (!shared) items.push_back((uint32_t)m | ((uint32_t)r << 12) | ((uint32_t)c << kItemCornerShift) | kItemSameObj);
                    }
                }
                continue;
            }
            uint32_t code = (uint32_t)a | ((uint32_t)b << 12);
            if (same) {
                code |= kCodeSameObj;
                const int dx = std::abs(leaves[a].gx - leaves[b].gx), dy = std::abs(leaves[a].gy - leaves[b].gy);
                if (objrecs[leaf_obj_slot[a]].is_robot && !(dx <= 1 && dy <= 1)) code |= kCodeCount;   // PhysicsEngine.cpp:856-863
                code |= (uint32_t)leaf_obj_slot[a] << kCodeObjShift;
            }
            pairs.push_back(code);
        }
    // per-edge flag: side of a CELL_RIGID voxel (visited by the ordered rigid-cell pass)
    std::vector<uint8_t> eflag(Ed, 0);
    for (int o = 0; o < n_obj; o++) {
        if (!dyn[o]) continue;
        for (const Boxel& b : W.objects[o].boxels)
            if (b.cell_type == CELL_RIGID)
                for (int q = 0; q < 4; q++) eflag[e.edge_local[b.ed[q]]] = 1;
    }
    // rigid-cell visit schedule: reference order = object, boxel, (top,right,bot,left); levels = point-disjoint
    std::vector<uint16_t> rlev(1, 0), rvis;
    std::vector<int> visit_order;      // dynamic-local edge id of every rigid-cell visit, reference order
    {
        std::vector<int> visit_edge, visit_level, last_level(Pd, 0);
        int n_levels = 0;
        for (int o = 0; o < n_obj; o++) {
            if (!dyn[o]) continue;
            for (const Boxel& b : W.objects[o].boxels) {
                if (b.cell_type != CELL_RIGID) continue;
                for (int q = 0; q < 4; q++) {
                    const int le = e.edge_local[b.ed[q]];
                    const int lv = std::max(last_level[ea[le]], last_level[eb[le]]) + 1;
                    last_level[ea[le]] = last_level[eb[le]] = lv;
                    visit_edge.push_back(le); visit_level.push_back(lv);
                    n_levels = std::max(n_levels, lv);
                }
            }
        }
        for (int lv = 1; lv <= n_levels; lv++) {
            for (size_t v = 0; v < visit_edge.size(); v++)
                if (visit_level[v] == lv) rvis.push_back((uint16_t)visit_edge[v]);
            rlev.push_back((uint16_t)rvis.size());
        }
        H.n_rlev = n_levels;
        H.n_rvis = (int)rvis.size();
        visit_order = visit_edge;
    }
    H.n_leaf = (int)leaves.size();
    H.n_obj = (int)objrecs.size();
    H.n_pairs = (int)pairs.size();
    H.n_items = (int)items.size();
    // group kernel: 128 candidates / survivors per warp and round (both buffers are drained in rounds, 64 at a time, so a
    // round keeps going after its first hits); the neighbour list holds
    // the whole pair table when that is small, else room for ~24 near neighbours per leaf (an overflowing warp falls back to
    // scanning its share of the table)
    const int n_warps = kGroupSizes[group_class_for(Pd, Ed)] / 32;
    H.cand_cap = 128 * n_warps;
    {
        const int want = std::min(H.n_pairs, std::max(2048, 24 * H.n_leaf));
        const int per_warp = ((want + n_warps - 1) / n_warps + 63) & ~63;
        H.list_cap = H.n_pairs ? per_warp * n_warps : 0;
        // items: a warp lists at most its own 32-wide chunks of the item table (chunks warp, warp + n_warps, ...)
        const int chunks = (H.n_items + 31) / 32;
        H.ilist_cap = ((chunks + n_warps - 1) / n_warps) * 32 * n_warps;
    }

    H.o_el0 = dw.begin_section(); dw.put_array(el0);
    H.o_eact = dw.begin_section(); dw.put_array(eact);
    H.o_fixed = dw.begin_section(); dw.put_array(fixed);
    H.o_leaf = dw.begin_section(); dw.put_array(leaves);
    H.o_obj = dw.begin_section(); dw.put_array(objrecs);
    H.o_pairs = dw.begin_section(); dw.put_array(pairs);
    H.o_rlev = dw.begin_section(); dw.put_array(rlev);
    H.o_rvis = dw.begin_section(); dw.put_array(rvis);
    H.o_items = dw.begin_section(); dw.put_array(items);
    // compact encodings: distinct (spring constant, init_length_eq) pairs + one packed word per edge
    bool warp_ok = Pd <= 255 && Ed <= 32767 && leaves.size() <= 32 && objrecs.size() <= 8 && visit_order.size() <= 255 * 4;
    std::vector<double> etab;                       // (ek, el0) pairs
    std::vector<uint32_t> eidx(Ed), eidx10(Ed);
    for (int k = 0; k < Ed; k++) {
        int type = -1;
        for (size_t q = 0; q + 1 < etab.size(); q += 2)
            if (etab[q] == ek[k] && etab[q + 1] == el0[k]) { type = (int)q / 2; break; }
        if (type < 0) { type = (int)etab.size() / 2; etab.push_back(ek[k]); etab.push_back(el0[k]); }
        if (type > 15) warp_ok = false;
        eidx[k] = (uint32_t)(ea[k] & 0xff) | ((uint32_t)(eb[k] & 0xff) << 8) | ((uint32_t)(type & 15) << kEidxTypeShift) |
                  ((uint32_t)eflag[k] << kEidxRigidBit);
        eidx10[k] = (uint32_t)(ea[k] & 0x3ff) | ((uint32_t)(eb[k] & 0x3ff) << kEidxBShift) | ((uint32_t)(type & 63) << kEidxTypeShift10) |
                    ((uint32_t)eflag[k] << kEidxRigidBit10);
    }
    H.n_types = (int)etab.size() / 2;
    if (H.n_types > kGroupMaxTypes) throw std::runtime_error("more than 64 distinct (spring constant, rest length) pairs in one environment");
    std::vector<uint16_t> pairs16, items16, lpt16;
    std::vector<uint8_t> rord8;
    if (warp_ok) {
        for (uint32_t c : pairs) {
            uint32_t v = (c & 31u) | (((c >> 12) & 31u) << 5);
            if (c & kCodeSameObj) v |= kPair16Same;
            if (c & kCodeCount) v |= kPair16Count | ((c >> kCodeObjShift) << kPair16ObjShift);
            pairs16.push_back((uint16_t)v);
        }
        for (uint32_t c : items)
            items16.push_back((uint16_t)((c & 31u) | (((c >> 12) & 31u) << 5) | (((c >> kItemCornerShift) & 3u) << 10) | kItem16Same));
        for (int le : visit_order) rord8.push_back((uint8_t)le);
        if (Ed > 255 && !visit_order.empty()) warp_ok = false;
    }
    for (const LeafRec& lr : leaves)
        for (int q = 0; q < 4; q++) lpt16.push_back(lr.pt[q]);
    std::vector<uint32_t> lpt8;
    H.fixed_lo = H.fixed_hi = 0;
    if (warp_ok) {
        for (const LeafRec& lr : leaves)
            lpt8.push_back((uint32_t)lr.pt[0] | ((uint32_t)lr.pt[1] << 8) | ((uint32_t)lr.pt[2] << 16) | ((uint32_t)lr.pt[3] << 24));
        for (int i = 0; i < Pd && i < 64; i++)
            if (fixed[i]) (i < 32 ? H.fixed_lo : H.fixed_hi) |= 1u << (i & 31);
    }
    H.warp_ok = warp_ok ? 1 : 0;
    H.o_csr8 = dw.begin_section(); dw.put_array(csr8);
    H.o_csrx = 0;
    if (!csrx.empty()) { H.o_csrx = dw.begin_section(); dw.put_array(csrx); }
    H.o_eidx = dw.begin_section(); dw.put_array(eidx);
    H.o_etab = dw.begin_section(); dw.put_array(etab);
    H.o_pairs16 = dw.begin_section(); dw.put_array(pairs16);
    H.o_items16 = dw.begin_section(); dw.put_array(items16);
    H.o_rord8 = dw.begin_section(); dw.put_array(rord8);
    H.o_eidx10 = dw.begin_section(); dw.put_array(eidx10);
    H.o_lpt16 = dw.begin_section(); dw.put_array(lpt16);
    H.o_lpt8 = dw.begin_section(); dw.put_array(lpt8);
    H.total = dw.begin_section();
    for (int q = 0; q < 16; q++) dw.put<uint8_t>(0);               // bulk copies round their size up to 16 bytes
    std::memcpy(design.data(), &H, sizeof(H));

    // ---- static blob
    statics.clear();
    if (!static_points.empty()) {
        BlobWriter sw(statics);
        StaticHeader S;
        std::memset(&S, 0, sizeof(S));
        statics.resize(sizeof(StaticHeader));
        S.Ps = (int)static_points.size();
        std::vector<double> spos(2 * static_points.size());
        for (size_t i = 0; i < static_points.size(); i++) {
            spos[2 * i] = e.pos[2 * static_points[i]];
            spos[2 * i + 1] = e.pos[2 * static_points[i] + 1];
        }
        std::vector<LeafRec> sleaves;
        std::vector<double> saabb;
        std::vector<int> cellx, celly;
        for (int o = 0; o < n_obj; o++) {
            if (dyn[o]) continue;
            const VoxelObject& ob = W.objects[o];
            for (int bi : ob.surface_boxels) {
                const Boxel& b = ob.boxels[bi];
                sleaves.push_back(make_leaf(ob, b, false, 0));
                double minx = 0, miny = 0, maxx = 0, maxy = 0;
                for (int q = 0; q < 4; q++) {
                    const double px = e.pos[2 * b.pt[q]], py = e.pos[2 * b.pt[q] + 1];
                    if (q == 0) { minx = maxx = px; miny = maxy = py; }
                    else { minx = std::fmin(minx, px); maxx = std::fmax(maxx, px); miny = std::fmin(miny, py); maxy = std::fmax(maxy, py); }
                }
                saabb.insert(saabb.end(), {minx, miny, maxx, maxy});
                // cell of the box centre on the voxel lattice (robust to object translations)
                cellx.push_back((int)std::floor(0.5 * (minx + maxx) / 0.1));
                celly.push_back((int)std::floor(0.5 * (miny + maxy) / 0.1));
            }
        }
        if ((int)sleaves.size() > kMaxLeaves) throw std::runtime_error("too many static surface voxels (max 4095)");
        S.n_leaf = (int)sleaves.size();
        std::vector<int16_t> grid;
        if (!sleaves.empty()) {
            const int x0 = *std::min_element(cellx.begin(), cellx.end()), x1 = *std::max_element(cellx.begin(), cellx.end());
            const int y0 = *std::min_element(celly.begin(), celly.end()), y1 = *std::max_element(celly.begin(), celly.end());
            S.gx0 = x0; S.gy0 = y0; S.gw = x1 - x0 + 1; S.gh = y1 - y0 + 1;
            grid.assign((size_t)S.gw * S.gh, (int16_t)-1);
            for (size_t l = 0; l < sleaves.size(); l++) {
                int16_t& slot = grid[(size_t)(celly[l] - y0) * S.gw + (cellx[l] - x0)];
                if (slot >= 0) throw std::runtime_error("two static voxels share one lattice cell");
                slot = (int16_t)l;
            }
        }
        S.o_pos = sw.begin_section(); sw.put_array(spos);
        S.o_leaf = sw.begin_section(); sw.put_array(sleaves);
        S.o_aabb = sw.begin_section(); sw.put_array(saabb);
        S.o_grid = sw.begin_section(); sw.put_array(grid);
        S.total = sw.begin_section();
        std::memcpy(statics.data(), &S, sizeof(S));
    }
}

void Batch::finalize() {
    CUDA_CHECK(cudaSetDevice(device_));
    for (void*& g : graphs_) if (g) { cudaGraphExecDestroy((cudaGraphExec_t)g); g = nullptr; }   // captured pointers go stale
    if (finalized_) download_all();
    const int n = num_envs();
    if (n == 0) throw std::runtime_error("finalize: no environments");

    std::vector<uint8_t> all_design, all_static;
    std::unordered_map<uint64_t, std::vector<std::pair<size_t, size_t>>> design_index, static_index;  // hash -> (offset, size)
    impl_->desc.assign(n, EnvDesc());
    impl_->class_envs.assign(kNumClasses, {});
    std::memset(impl_->class_max, 0, sizeof(impl_->class_max));
    impl_->class_smem.assign(kNumClasses, 0);
    impl_->env_cost.clear();
    size_t state_doubles = 0, scratch_bytes = 0;
    std::memset(impl_->class_ppt, 0, sizeof(impl_->class_ppt));
    std::memset(impl_->class_xd, 0, sizeof(impl_->class_xd));
    std::memset(impl_->class_slots, 0, sizeof(impl_->class_slots));
    act_stride_ = 1;

    auto intern = [](std::vector<uint8_t>& pool, std::unordered_map<uint64_t, std::vector<std::pair<size_t, size_t>>>& index,
                     const std::vector<uint8_t>& blob) -> size_t {
        const uint64_t h = fnv1a(blob);
        for (auto& cand : index[h])
            if (cand.second == blob.size() && std::memcmp(pool.data() + cand.first, blob.data(), blob.size()) == 0) return cand.first;
        const size_t off = pool.size();
        pool.insert(pool.end(), blob.begin(), blob.end());
        index[h].push_back({off, blob.size()});
        return off;
    };

    // Packing one environment (pair tables, gather lists, static grid) touches only that environment: done on all host
    // cores, a chunk at a time; interning, offsets and class lists follow sequentially in environment order.  This is the
    // rebuild path of design loops (`env.update(body, connections)`, evogym/envs/walk.py:105-120), where thousands of robots
    // change between episodes.
    constexpr int kChunk = 2048;
    const int n_threads = (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    std::vector<std::vector<uint8_t>> chunk_design(std::min(n, kChunk)), chunk_static(std::min(n, kChunk));
    for (int i = 0; i < n; i++) {
        EnvHost& e = envs_[i];
        if (i % kChunk == 0) {
            const int base = i, m = std::min(kChunk, n - base);
            for (int k = 0; k < m; k++)
                if (envs_[base + k].world.objects.empty())
                    throw std::runtime_error("finalize: environment " + std::to_string(base + k) + " has no objects");
            std::atomic<int> next(0);
            std::vector<std::exception_ptr> errors(n_threads);
            auto worker = [&](int tid) {
                try {
                    for (int k = next.fetch_add(1); k < m; k = next.fetch_add(1)) pack_env(base + k, chunk_design[k], chunk_static[k]);
                } catch (...) { errors[tid] = std::current_exception(); }
            };
            if (n_threads == 1 || m < 8) worker(0);
            else {
                std::vector<std::thread> pool;
                for (int t = 0; t < n_threads; t++) pool.emplace_back(worker, t);
                for (auto& th : pool) th.join();
            }
            for (auto& ep : errors) if (ep) std::rethrow_exception(ep);
        }
        const std::vector<uint8_t>& design = chunk_design[i % kChunk];
        const std::vector<uint8_t>& statics = chunk_static[i % kChunk];
        const DesignHeader* H = reinterpret_cast<const DesignHeader*>(design.data());
        EnvDesc& d = impl_->desc[i];
        d.design_off = intern(all_design, design_index, design);
        d.static_off = statics.empty() ? ~0ull : intern(all_static, static_index, statics);
        d.state_off = state_doubles;
        e.state_off = state_doubles;
        size_t sd = 8 * (size_t)H->Pd + (size_t)H->Ed;
        sd = (sd + 1) & ~(size_t)1;
        state_doubles += sd;
        act_stride_ = std::max(act_stride_, e.n_act);
        e.cls = class_for(*H, impl_->use_warp_kernel);
        {   // development aid (MEVO_DUMP_CLASSES=1): largest table sizes per size class, printed at the end of finalize
            int* m = impl_->class_max[e.cls];
            const int v[6] = {H->Pd, H->Ed, H->n_leaf, H->n_pairs, H->n_items, H->n_rvis};
            for (int q = 0; q < 6; q++) m[q] = std::max(m[q], v[q]);
        }
        if (e.cls >= kFirstGroupClass) {
            const GLayout L = g_layout(H->Pd, H->Ed, H->n_leaf, H->n_types, H->cand_cap, H->n_rvis, H->n_rlev);
            if (L.total > kMaxSmemPerBlock) throw std::runtime_error("environment " + std::to_string(i) + " does not fit in shared memory");
            impl_->class_smem[e.cls] = std::max(impl_->class_smem[e.cls], L.total);
            impl_->class_ppt[e.cls] = std::max(impl_->class_ppt[e.cls], (H->Pd + kGroupSizes[e.cls] - 1) / kGroupSizes[e.cls]);
            if (impl_->class_ppt[e.cls] == 3) impl_->class_ppt[e.cls] = 4;
            if (H->o_csrx) impl_->class_xd[e.cls] = kPptExtBit;
            d.scratch_off16 = (int)(scratch_bytes / 16);
            scratch_bytes += ((size_t)(H->list_cap + H->ilist_cap) * 4 + 15) & ~(size_t)15;
            if (scratch_bytes / 16 > 0x7fffffffull) throw std::runtime_error("neighbour-list scratch exceeds 32 GB");
        }
        impl_->class_envs[e.cls].push_back(i);
        // work estimate of one substep (spring passes, point passes, corner items): environments of similar cost
        // are placed in the same CTA so that the per-substep CTA barrier waits as little as possible
        impl_->env_cost.push_back(1000 * ((H->Pd + 31) / 32) + 700 * ((H->Ed + 63) / 64) + 40 * ((H->n_items + 31) / 32) +
                                  40 * ((H->n_pairs + 31) / 32) + 20 * H->n_rvis);
    }
    if (const char* e = std::getenv("MEVO_DUMP_CLASSES"))
        if (std::atoi(e))
            for (int c = 0; c < kNumClasses; c++)
                if (!impl_->class_envs[c].empty())
                    std::fprintf(stderr, "[mevo] class %d (%d threads, %d points per thread): %zu envs, max points %d edges %d leaves %d pairs %d "
                                 "items %d visits %d, shared memory %d B per CTA\n", c, kGroupSizes[c], impl_->class_ppt[c],
                                 impl_->class_envs[c].size(), impl_->class_max[c][0], impl_->class_max[c][1], impl_->class_max[c][2],
                                 impl_->class_max[c][3], impl_->class_max[c][4], impl_->class_max[c][5], impl_->class_smem[c]);
    for (int c = 0; c < kFirstGroupClass; c++)
        std::stable_sort(impl_->class_envs[c].begin(), impl_->class_envs[c].end(),
                         [&](int a, int b) { return impl_->env_cost[a] > impl_->env_cost[b]; });
    for (int i = 0; i < n; i++) impl_->desc[i].act_off = i * act_stride_;

    // (re)allocate
    cudaFree(d_design_); cudaFree(d_static_); cudaFree(d_state_); cudaFree(d_state0_); cudaFree(d_act_);
    cudaFree(d_envdesc_); cudaFree(d_class_envs_); cudaFree(d_unstable_); cudaFree(d_time_); cudaFree(d_err_);
    d_design_ = d_static_ = nullptr; d_state_ = d_state0_ = d_act_ = nullptr;
    state_doubles_ = state_doubles;
    CUDA_CHECK(cudaMalloc(&d_design_, all_design.size() + 16));
    CUDA_CHECK(cudaMalloc(&d_static_, all_static.size() + 16));
    CUDA_CHECK(cudaMalloc(&d_state_, sizeof(double) * (state_doubles + 2)));
    CUDA_CHECK(cudaMalloc(&d_state0_, sizeof(double) * (state_doubles + 2)));
    CUDA_CHECK(cudaMalloc(&d_act_, sizeof(double) * (size_t)n * act_stride_));
    CUDA_CHECK(cudaMalloc(&d_envdesc_, sizeof(EnvDesc) * n));
    CUDA_CHECK(cudaMalloc(&d_class_envs_, sizeof(int) * n));
    CUDA_CHECK(cudaMalloc(&d_unstable_, n));
    CUDA_CHECK(cudaMalloc(&d_time_, sizeof(int) * n));
    CUDA_CHECK(cudaMalloc(&d_err_, sizeof(int) * n));
    cudaFree(impl_->d_scratch); impl_->d_scratch = nullptr;
    cudaFree(impl_->d_cf); impl_->d_cf = nullptr;
    for (int k = 0; k < Impl::kSlots; k++) {        // snapshot slots hold the old packing: dropped
        cudaFree(impl_->d_slot[k]); impl_->d_slot[k] = nullptr;
        cudaFree(impl_->d_slot_time[k]); impl_->d_slot_time[k] = nullptr;
    }
    CUDA_CHECK(cudaMalloc(&impl_->d_scratch, scratch_bytes + 16));
    impl_->scratch_bytes = scratch_bytes;
    {   // item counters of the chunked warp-kernel launches: 1 + one word per CTA-group, per warp class
        int words = 0;
        for (int c = 0; c < kFirstGroupClass; c++) {
            impl_->work_off[c] = words;
            words += 4 + (int)impl_->class_envs[c].size() / 4;           // groups are >= 4 environments wide
        }
        cudaFree(impl_->d_work); impl_->d_work = nullptr;
        CUDA_CHECK(cudaMalloc(&impl_->d_work, sizeof(int) * (size_t)words));
    }
    // dynamic shared-memory limits of the kernels this batch launches, on this batch's device
    for (int c = 0; c < kNumClasses; c++) {
        if (impl_->class_envs[c].empty()) continue;
        if (c == 0) {
            if (impl_->wepc == 8) configure_wstep<MEVO_WCAPS_A, 8, 2>();
            else if (impl_->wepc == 4) configure_wstep<MEVO_WCAPS_A, 4, 4>();
            else configure_wstep<MEVO_WCAPS_A, 16, 1>();
        } else if (c == 1) configure_wstep<MEVO_WCAPS_C, 12, 1>();
        else if (c == 2) configure_wstep<MEVO_WCAPS_B, 10, 1>();
        else {
            gstep_dispatch(c, impl_->class_ppt[c] | impl_->class_xd[c], GAction::Configure, nullptr, impl_->class_smem[c], nullptr);
            // resident CTAs: the register budget of the build (64 / 128 / 256 / 512 threads -> 8 / 4 / 2 / 1 per SM) or shared memory
            const int by_regs = 512 / kGroupSizes[c], by_smem = (228 * 1024) / (impl_->class_smem[c] + 1024);
            impl_->class_slots[c] = impl_->n_sm * std::max(1, std::min(by_regs, by_smem));
        }
    }
    if (!d_events_) { CUDA_CHECK(cudaMalloc(&d_events_, sizeof(unsigned long long))); CUDA_CHECK(cudaMemset(d_events_, 0, sizeof(unsigned long long))); }
    d_act_ext_ = nullptr;
    device_bytes_ = all_design.size() + all_static.size() + scratch_bytes + 2 * sizeof(double) * state_doubles +
                    sizeof(double) * (size_t)n * act_stride_ + sizeof(EnvDesc) * n + 10 * (size_t)n;
    CUDA_CHECK(cudaMemcpy(d_design_, all_design.data(), all_design.size(), cudaMemcpyHostToDevice));
    if (!all_static.empty()) CUDA_CHECK(cudaMemcpy(d_static_, all_static.data(), all_static.size(), cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(d_envdesc_, impl_->desc.data(), sizeof(EnvDesc) * n, cudaMemcpyHostToDevice));
    {
        std::vector<int> flat;
        impl_->class_off.assign(kNumClasses, 0);
        for (int c = 0; c < kNumClasses; c++) {
            impl_->class_off[c] = flat.size();
            flat.insert(flat.end(), impl_->class_envs[c].begin(), impl_->class_envs[c].end());
        }
        CUDA_CHECK(cudaMemcpy(d_class_envs_, flat.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    }
    {
        std::vector<double> ones((size_t)n * act_stride_, 1.0);   // action 1.0 == rest length
        CUDA_CHECK(cudaMemcpy(d_act_, ones.data(), sizeof(double) * ones.size(), cudaMemcpyHostToDevice));
        std::vector<int> times(n);
        for (int i = 0; i < n; i++) times[i] = (int)envs_[i].sim_time;
        CUDA_CHECK(cudaMemcpy(d_time_, times.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemset(d_unstable_, 0, n));
        CUDA_CHECK(cudaMemset(d_err_, 0, sizeof(int) * n));
    }
    if (impl_->unstable_host) cudaFreeHost(impl_->unstable_host);
    impl_->unstable_host = nullptr;
    CUDA_CHECK(cudaMallocHost(&impl_->unstable_host, (size_t)std::max(n, 1)));
    std::memset(impl_->unstable_host, 0, (size_t)std::max(n, 1));
    impl_->time_host.assign(n, 0);
    impl_->com_obj = -1;
    if (cand_cap_ > 0) enable_candidate_dump(cand_cap_);
    if (force_dump_) enable_force_dump();

    // state + reset snapshot
    const bool first = impl_->pos0.empty();
    if (first) { impl_->pos0.resize(n); impl_->vel0.resize(n); }
    impl_->pos0.resize(n); impl_->vel0.resize(n);
    std::vector<double> host_state(state_doubles + 2, 0.0), host_state0(state_doubles + 2, 0.0);
    for (int i = 0; i < n; i++) {
        EnvHost& e = envs_[i];
        if (impl_->pos0[i].size() != e.pos.size()) { impl_->pos0[i] = e.pos; impl_->vel0[i] = e.vel; }
        const int Pd = (int)e.dyn_points.size(), Ed = (int)e.dyn_edges.size();
        double* s = host_state.data() + e.state_off;
        double* s0 = host_state0.data() + e.state_off;
        for (int k = 0; k < Pd; k++) {
            const int g = e.dyn_points[k];
            for (int c = 0; c < 2; c++) {
                s[2 * k + c] = e.pos[2 * g + c];
                s[2 * Pd + 2 * k + c] = e.vel[2 * g + c];
                s[4 * Pd + 2 * k + c] = e.pos_last[2 * g + c];
                s[6 * Pd + 2 * k + c] = e.vel_true[2 * g + c];
                s0[2 * k + c] = impl_->pos0[i][2 * g + c];
                s0[2 * Pd + 2 * k + c] = impl_->vel0[i][2 * g + c];
            }
        }
        for (int k = 0; k < Ed; k++) s[8 * Pd + k] = e.goal[e.dyn_edges[k]];
        e.host_fresh = true;
    }
    CUDA_CHECK(cudaMemcpy(d_state_, host_state.data(), sizeof(double) * (state_doubles + 2), cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(d_state0_, host_state0.data(), sizeof(double) * (state_doubles + 2), cudaMemcpyHostToDevice));
    finalized_ = true;
    finalize_epoch_++;
    impl_->com_obj = -1;              // per-object point ranges are rebuilt on the next use
}

// ------------------------------------------------------------------------------------------------
// state movement

void Batch::download_state(int env) {
    EnvHost& e = envs_.at(env);
    if (!finalized_ || e.host_fresh) return;
    CUDA_CHECK(cudaSetDevice(device_));
    const int Pd = (int)e.dyn_points.size(), Ed = (int)e.dyn_edges.size();
    std::vector<double> s(8 * (size_t)Pd + Ed);
    CUDA_CHECK(cudaMemcpyAsync(s.data(), d_state_ + e.state_off, sizeof(double) * s.size(), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    int t = 0;
    CUDA_CHECK(cudaMemcpyAsync(&t, d_time_ + env, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    for (int k = 0; k < Pd; k++) {
        const int g = e.dyn_points[k];
        for (int c = 0; c < 2; c++) {
            e.pos[2 * g + c] = s[2 * k + c];
            e.vel[2 * g + c] = s[2 * Pd + 2 * k + c];
            e.pos_last[2 * g + c] = s[4 * Pd + 2 * k + c];
            e.vel_true[2 * g + c] = s[6 * Pd + 2 * k + c];
        }
    }
    for (int k = 0; k < Ed; k++) e.goal[e.dyn_edges[k]] = s[8 * Pd + k];
    e.sim_time = t;
    e.host_fresh = true;
}

void Batch::download_all() {
    if (!finalized_) return;
    CUDA_CHECK(cudaSetDevice(device_));
    bool any = false;
    for (auto& e : envs_) any |= !e.host_fresh;
    if (!any) return;
    std::vector<double> all(state_doubles_ + 2);
    std::vector<int> times(num_envs());
    CUDA_CHECK(cudaMemcpyAsync(all.data(), d_state_, sizeof(double) * state_doubles_, cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaMemcpyAsync(times.data(), d_time_, sizeof(int) * num_envs(), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    for (int i = 0; i < num_envs(); i++) {
        EnvHost& e = envs_[i];
        if (e.host_fresh) continue;
        const int Pd = (int)e.dyn_points.size(), Ed = (int)e.dyn_edges.size();
        const double* s = all.data() + e.state_off;
        for (int k = 0; k < Pd; k++) {
            const int g = e.dyn_points[k];
            for (int c = 0; c < 2; c++) {
                e.pos[2 * g + c] = s[2 * k + c];
                e.vel[2 * g + c] = s[2 * Pd + 2 * k + c];
                e.pos_last[2 * g + c] = s[4 * Pd + 2 * k + c];
                e.vel_true[2 * g + c] = s[6 * Pd + 2 * k + c];
            }
        }
        for (int k = 0; k < Ed; k++) e.goal[e.dyn_edges[k]] = s[8 * Pd + k];
        e.sim_time = times[i];
        e.host_fresh = true;
    }
}

void Batch::upload_state(int env) {
    EnvHost& e = envs_.at(env);
    CUDA_CHECK(cudaSetDevice(device_));
    const int Pd = (int)e.dyn_points.size(), Ed = (int)e.dyn_edges.size();
    std::vector<double> s(8 * (size_t)Pd + Ed);
    for (int k = 0; k < Pd; k++) {
        const int g = e.dyn_points[k];
        for (int c = 0; c < 2; c++) {
            s[2 * k + c] = e.pos[2 * g + c];
            s[2 * Pd + 2 * k + c] = e.vel[2 * g + c];
            s[4 * Pd + 2 * k + c] = e.pos_last[2 * g + c];
            s[6 * Pd + 2 * k + c] = e.vel_true[2 * g + c];
        }
    }
    for (int k = 0; k < Ed; k++) s[8 * Pd + k] = e.goal[e.dyn_edges[k]];
    const int t = (int)e.sim_time;
    CUDA_CHECK(cudaMemcpyAsync(d_state_ + e.state_off, s.data(), sizeof(double) * s.size(), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    CUDA_CHECK(cudaMemcpyAsync(d_time_ + env, &t, sizeof(int), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    e.host_fresh = true;
}

void Batch::get_state(int env, double* pos, double* vel, double* pos_last, double* vel_true) {
    download_state(env);
    const EnvHost& e = envs_.at(env);
    const size_t bytes = sizeof(double) * e.pos.size();
    if (pos) std::memcpy(pos, e.pos.data(), bytes);
    if (vel) std::memcpy(vel, e.vel.data(), bytes);
    if (pos_last) std::memcpy(pos_last, e.pos_last.data(), bytes);
    if (vel_true) std::memcpy(vel_true, e.vel_true.data(), bytes);
}

void Batch::set_state(int env, const double* pos, const double* vel, const double* pos_last, const double* vel_true) {
    download_state(env);
    EnvHost& e = envs_.at(env);
    // static (all-fixed, pruned) points keep their own values: they cannot move
    for (size_t g = 0; g < e.pos.size() / 2; g++) {
        if (finalized_ && e.point_local[g] < 0) continue;
        for (int c = 0; c < 2; c++) {
            if (pos) e.pos[2 * g + c] = pos[2 * g + c];
            if (vel) e.vel[2 * g + c] = vel[2 * g + c];
            if (pos_last) e.pos_last[2 * g + c] = pos_last[2 * g + c];
            if (vel_true) e.vel_true[2 * g + c] = vel_true[2 * g + c];
        }
    }
    if (finalized_) upload_state(env);
}

void Batch::get_goal(int env, double* goal) {
    download_state(env);
    const EnvHost& e = envs_.at(env);
    std::memcpy(goal, e.goal.data(), sizeof(double) * e.goal.size());
}

void Batch::set_goal(int env, const double* goal) {
    download_state(env);
    EnvHost& e = envs_.at(env);
    std::memcpy(e.goal.data(), goal, sizeof(double) * e.goal.size());
    if (finalized_) upload_state(env);
}

long Batch::get_time(int env) {
    download_state(env);
    return envs_.at(env).sim_time;
}

void Batch::set_time(int env, long t) {
    download_state(env);
    envs_.at(env).sim_time = t;
    if (finalized_) upload_state(env);
}

void Batch::translate_object(int env, const std::string& name, double dx, double dy) {
    // Environment::translate_object (Environment.cpp:381-392): only points_pos moves; pos_last stays, so the
    // next substep sees a true-velocity spike -- reproduced by simulating the object from now on.
    EnvHost& e = envs_.at(env);
    const int o = e.world.find_object(name);
    if (o < 0) return;
    if (finalized_) download_all();
    const VoxelObject& ob = e.world.objects[o];
    for (int p = ob.min_point; p <= ob.max_point; p++) { e.pos[2 * p] += dx; e.pos[2 * p + 1] += dy; }
    const bool was_static = ob.all_fixed && !ob.is_robot && !e.obj_force_dynamic[o];
    if (was_static) {
        e.obj_force_dynamic[o] = 1;
        if (finalized_) { finalized_ = false; finalize(); }
    } else if (finalized_) {
        upload_state(env);
    }
}

// ------------------------------------------------------------------------------------------------
// actions

void Batch::set_actions(const double* host_values, int stride) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) throw std::runtime_error("set_actions before finalize");
    if (stride == act_stride_) {
        CUDA_CHECK(cudaMemcpyAsync(d_act_, host_values, sizeof(double) * (size_t)num_envs() * stride, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    } else {
        CUDA_CHECK(cudaMemcpy2DAsync(d_act_, sizeof(double) * act_stride_, host_values, sizeof(double) * stride,
                                     sizeof(double) * std::min(stride, act_stride_), num_envs(), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    }
}

void Batch::set_actions_device(const double* dev_values, int stride) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) throw std::runtime_error("set_actions before finalize");
    CUDA_CHECK(cudaMemcpy2DAsync(d_act_, sizeof(double) * act_stride_, dev_values, sizeof(double) * stride,
                                 sizeof(double) * std::min(stride, act_stride_), num_envs(), cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
}

void Batch::bind_actions_device(const double* dev_values, int stride) {
    if (dev_values && stride < act_stride_) throw std::runtime_error("bind_actions_device: stride smaller than the batch's action stride");
    d_act_ext_ = dev_values;
    act_ext_stride_ = stride;
}

unsigned long long Batch::unstable_events() {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!d_events_) return 0;
    sync();
    unsigned long long v = 0;
    CUDA_CHECK(cudaMemcpy(&v, d_events_, sizeof(v), cudaMemcpyDeviceToHost));
    return v;
}

bool Batch::env_set_action(int env, const std::string& robot, const double* action, int n) {
    // Environment::set_robot_action + PhysicsEngine::update_actuator_goals (Environment.cpp:207-237,
    // PhysicsEngine.cpp:38-97), evaluated on the host mirror of length_eq_goal, then uploaded.
    EnvHost& e = envs_.at(env);
    const int o = e.world.find_object(robot);
    if (o < 0) { std::printf("Error: No robot named %s.\n", robot.c_str()); return false; }
    const VoxelObject& ob = e.world.objects[o];
    if (!ob.is_robot) { std::printf("Error: %s is not robot.\n", robot.c_str()); return false; }
    std::vector<const Boxel*> acts;
    for (int i = 0; i < n; i++) {
        auto it = ob.actuator_grid_to_boxel.find((int)action[2 * i]);
        if (it == ob.actuator_grid_to_boxel.end()) {
            // the reference does `throw;` here -> std::terminate (Robot.cpp:28-35); we refuse instead
            std::printf("Error: Robot action inproperly defined - no actuator found at index.\n");
            return false;
        }
        acts.push_back(&ob.boxels[it->second]);
    }
    download_state(env);
    std::map<int, int> count;
    auto driven = [](const Boxel* b, int& e0, int& e1) {
        if (b->cell_type == CELL_ACT_H) { e0 = b->ed[0]; e1 = b->ed[2]; }
        else { e0 = b->ed[3]; e1 = b->ed[1]; }
    };
    for (const Boxel* b : acts) { int e0, e1; driven(b, e0, e1); e.goal[e0] = 0; e.goal[e1] = 0; count[e0] = 0; count[e1] = 0; }
    for (int i = 0; i < n; i++) {
        int e0, e1; driven(acts[i], e0, e1);
        e.goal[e0] += action[2 * i + 1]; e.goal[e1] += action[2 * i + 1];
        count[e0] += 1; count[e1] += 1;
    }
    for (auto& kv : count) e.goal[kv.first] *= e.world.el0[kv.first] / kv.second;
    if (finalized_) upload_state(env);
    return true;
}

// ------------------------------------------------------------------------------------------------
// stepping

void Batch::step_async(int n_steps, bool apply_actions, int n_substeps) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    StepParams prm;
    prm.design = d_design_;
    prm.statics = d_static_;
    prm.state = d_state_;
    prm.act = d_act_ext_ ? d_act_ext_ : d_act_;
    prm.act_stride = d_act_ext_ ? act_ext_stride_ : act_stride_;
    prm.state0 = d_state0_;
    prm.auto_reset = auto_reset_ ? 1 : 0;
    prm.unstable_events = d_events_;
    prm.envs = reinterpret_cast<const EnvDesc*>(d_envdesc_);
    prm.n_steps = n_steps;
    prm.n_substeps = n_substeps;
    prm.apply_actions = apply_actions ? 1 : 0;
    prm.cta_sync = impl_->cta_sync;
    prm.unstable = d_unstable_;
    prm.sim_time = d_time_;
    prm.err = d_err_;
    prm.dbg_cand = d_cand_;
    prm.dbg_cand_n = d_cand_n_;
    prm.dbg_cand_cap = cand_cap_;
    prm.dbg_cf = impl_->d_cf;
    prm.scratch = impl_->d_scratch;
    prm.chunk_first = prm.chunk_last = 1;
    prm.work = nullptr;
    prm.n_chunks = 1;
    cudaStream_t main_st = (cudaStream_t)stream_;
    stats_.launches = 0;
    int n_used = 0;
    int n_group = 0;
    double waves = 0.0;                       // CTA demand of the group classes over what the device holds at once
    for (int c = 0; c < kNumClasses; c++) {
        const bool used = !impl_->class_envs[c].empty();
        n_used += used ? 1 : 0;
        if (used && c >= kFirstGroupClass) { n_group++; waves += (double)impl_->class_envs[c].size() / std::max(1, impl_->class_slots[c]); }
    }
    // chunked, laned launches (see Impl::chunks).  One launch leaves about 1 - waves / ceil(waves) of the CTA slots idle; with n
    // chunks the figure is that of n * waves, and every extra launch costs its stage-in and a list build (~1.5 % of a step).
    // The chunk count with the best estimate is taken, if it gains anything: 1024 robots of 20x20 (3.46 waves) run as 2 chunks
    // (+10 %), the 8192-robot mix (15 waves) as one launch (chunked: -9 %).  Only a plain single env-step is chunked; the debug
    // dumps describe "the first substep of the launch" and keep one launch
    int n_chunks = 1;
    auto idle = [](double w) { return 1.0 - w / std::ceil(w - 1e-9); };
    const bool chunkable = n_steps == 1 && !d_cand_ && !impl_->d_cf;
    const bool forced = impl_->chunk_min_waves <= 0.0;                       // tests: the largest chunk count whatever the population
    if (chunkable && n_group >= 1 && impl_->chunks > 1) {
        double best = impl_->chunk_min_gain;
        for (int k = 2; k <= impl_->chunks && 2 * k <= n_substeps; k++) {
            const double gain = idle(waves) - idle(waves * k) - 0.015 * (k - 1);
            if (forced) n_chunks = k;
            else if (waves > impl_->chunk_min_waves && gain > best) { best = gain; n_chunks = k; }
        }
    }
    const bool laned = n_chunks > 1;
    const bool fork = (n_used > 1 && impl_->concurrent_classes != 0) || laned;
    if (fork) {
        if (!impl_->fork) CUDA_CHECK(cudaEventCreateWithFlags(&impl_->fork, cudaEventDisableTiming));
        CUDA_CHECK(cudaEventRecord(impl_->fork, main_st));
    }
    struct Lane { int c, off, n; cudaStream_t st; cudaEvent_t done; };
    std::vector<Lane> lanes;
    auto launch_class = [&](int c, const StepParams& p, bool balance, cudaStream_t st) {
        switch (c) {
            case 0:
                // default: one 16-warp CTA per SM (all warps of an SM in the same code region)
                if (impl_->wepc == 8) launch_wstep<MEVO_WCAPS_A, 8, 2>(p, impl_->n_sm, balance, st);
                else if (impl_->wepc == 4) launch_wstep<MEVO_WCAPS_A, 4, 4>(p, impl_->n_sm, balance, st);
                else launch_wstep<MEVO_WCAPS_A, 16, 1>(p, impl_->n_sm, balance, st);
                break;
            case 1: launch_wstep<MEVO_WCAPS_C, 12, 1>(p, impl_->n_sm, balance, st); break;
            case 2: launch_wstep<MEVO_WCAPS_B, 10, 1>(p, impl_->n_sm, balance, st); break;
            default: gstep_dispatch(c, impl_->class_ppt[c] | impl_->class_xd[c], GAction::Launch, &p, impl_->class_smem[c], st); break;
        }
        CUDA_CHECK(cudaGetLastError());
        stats_.launches++;
    };
    for (int c = kNumClasses - 1; c >= 0; c--) {
        const int n = (int)impl_->class_envs[c].size();
        if (n == 0) continue;
        if (laned && c >= kFirstGroupClass) {
            const int L = std::max(1, std::min(impl_->lanes, n));
            for (int l = 0; l < L; l++) {
                if (!impl_->lane_stream[c][l]) {
                    CUDA_CHECK(cudaStreamCreateWithFlags(&impl_->lane_stream[c][l], cudaStreamNonBlocking));
                    CUDA_CHECK(cudaEventCreateWithFlags(&impl_->lane_done[c][l], cudaEventDisableTiming));
                }
                const int lo = (int)((long long)n * l / L), hi = (int)((long long)n * (l + 1) / L);
                CUDA_CHECK(cudaStreamWaitEvent(impl_->lane_stream[c][l], impl_->fork, 0));
                lanes.push_back({c, lo, hi - lo, impl_->lane_stream[c][l], impl_->lane_done[c][l]});
            }
            continue;
        }
        cudaStream_t st = main_st;
        if (fork) {
            if (!impl_->class_stream[c]) {
                CUDA_CHECK(cudaStreamCreateWithFlags(&impl_->class_stream[c], cudaStreamNonBlocking));
                CUDA_CHECK(cudaEventCreateWithFlags(&impl_->class_done[c], cudaEventDisableTiming));
            }
            st = impl_->class_stream[c];
            CUDA_CHECK(cudaStreamWaitEvent(st, impl_->fork, 0));
        }
        prm.class_envs = d_class_envs_ + impl_->class_off[c];
        prm.n_class_envs = n;
        prm.work = nullptr;
        prm.n_chunks = 1;
        if (c < kFirstGroupClass && chunkable && impl_->wchunks > 1) {
            // warp classes: one CTA per (CTA-group, chunk) item where that fills the SMs better than one CTA per group
            const int unit = c == 0 ? (impl_->wepc ? impl_->wepc : 16) : c == 1 ? 12 : 10;
            const int slots = impl_->n_sm * (c == 0 && impl_->wepc ? 16 / impl_->wepc : 1);
            const double w = (double)((n + unit - 1) / unit) / slots;
            double best = impl_->chunk_min_gain;
            for (int k = 2; k <= impl_->wchunks && 2 * k <= n_substeps; k++) {
                const double gain = idle(w) - idle(w * k) - 0.012 * (k - 1);
                if (forced) prm.n_chunks = k;
                else if ((w > 1.0 || n_used > 1) && gain > best) { best = gain; prm.n_chunks = k; }     // alone and resident at once: nothing to fill
            }
            if (prm.n_chunks > 1) prm.work = impl_->d_work + impl_->work_off[c];
        }
        launch_class(c, prm, impl_->balance != 0, st);
        prm.work = nullptr;
        prm.n_chunks = 1;
        if (fork) {
            CUDA_CHECK(cudaEventRecord(impl_->class_done[c], st));
            CUDA_CHECK(cudaStreamWaitEvent(main_st, impl_->class_done[c], 0));
        }
    }
    for (int k = 0; k < n_chunks && !lanes.empty(); k++) {
        const int s0 = (int)((long long)n_substeps * k / n_chunks), s1 = (int)((long long)n_substeps * (k + 1) / n_chunks);
        StepParams cp = prm;
        cp.n_substeps = s1 - s0;
        cp.chunk_first = k == 0;
        cp.chunk_last = k == n_chunks - 1;
        cp.apply_actions = (apply_actions && k == 0) ? 1 : 0;
        for (const Lane& ln : lanes) {
            cp.class_envs = d_class_envs_ + impl_->class_off[ln.c] + ln.off;
            cp.n_class_envs = ln.n;
            launch_class(ln.c, cp, false, ln.st);
        }
    }
    for (const Lane& ln : lanes) {
        CUDA_CHECK(cudaEventRecord(ln.done, ln.st));
        CUDA_CHECK(cudaStreamWaitEvent(main_st, ln.done, 0));
    }
    for (auto& e : envs_) e.host_fresh = false;
}

void Batch::sync() {
    CUDA_CHECK(cudaSetDevice(device_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

// Sim::step for every environment like step_async, without allocating anything: no cudaMalloc / cudaFree, no stream or event
// creation (the size classes run back to back on the batch's stream), neighbour lists in the file-scope g_static_scratch.
// Inputs and outputs are the caller's: `dev_actions` (device, values[env * stride + slot]; nullptr keeps the edge goals as last
// set) and `dev_unstable_out` (device, num_envs bytes; nullptr = the batch's own flag buffer).  The state stays in the batch.
void Batch::step_static_scratch(int n_steps, const double* dev_actions, int stride, uint8_t* dev_unstable_out) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_)
        throw std::runtime_error("step_static_scratch: call finalize first (finalize allocates; this entry point never does)");
    if (n_steps < 1) throw std::runtime_error("step_static_scratch: n_steps must be >= 1");
    if (impl_->scratch_bytes > kStaticScratchBytes)
        throw std::runtime_error("step_static_scratch: this batch needs " + std::to_string(impl_->scratch_bytes) +
                                 " bytes of neighbour-list scratch, the static array holds " + std::to_string(kStaticScratchBytes) +
                                 " (MEVO_STATIC_SCRATCH_BYTES): use fewer / smaller environments per batch or mevo_batch_step_async");
    if (dev_actions && stride < act_stride_)
        throw std::runtime_error("step_static_scratch: stride smaller than the batch's action stride");
    StepParams prm;
    prm.design = d_design_;
    prm.statics = d_static_;
    prm.state = d_state_;
    prm.act = dev_actions ? dev_actions : d_act_;
    prm.act_stride = dev_actions ? stride : act_stride_;
    prm.state0 = d_state0_;
    prm.auto_reset = auto_reset_ ? 1 : 0;
    prm.unstable_events = d_events_;
    prm.envs = reinterpret_cast<const EnvDesc*>(d_envdesc_);
    prm.n_steps = n_steps;
    prm.n_substeps = 30;
    prm.apply_actions = dev_actions ? 1 : 0;
    prm.cta_sync = impl_->cta_sync;
    prm.unstable = dev_unstable_out ? dev_unstable_out : d_unstable_;
    prm.sim_time = d_time_;
    prm.err = d_err_;
    prm.dbg_cand = nullptr;
    prm.dbg_cand_n = nullptr;
    prm.dbg_cand_cap = 0;
    prm.dbg_cf = nullptr;
    prm.chunk_first = prm.chunk_last = 1;
    prm.work = nullptr;
    prm.n_chunks = 1;
    prm.scratch = static_scratch_address(device_);          // NOT g_static_scratch itself: that name is a host-side stand-in
    cudaStream_t st = (cudaStream_t)stream_;
    stats_.launches = 0;
    for (int c = kNumClasses - 1; c >= 0; c--) {
        const int n = (int)impl_->class_envs[c].size();
        if (n == 0) continue;
        prm.class_envs = d_class_envs_ + impl_->class_off[c];
        prm.n_class_envs = n;
        switch (c) {
            case 0:
                if (impl_->wepc == 8) launch_wstep<MEVO_WCAPS_A, 8, 2>(prm, impl_->n_sm, impl_->balance != 0, st);
                else if (impl_->wepc == 4) launch_wstep<MEVO_WCAPS_A, 4, 4>(prm, impl_->n_sm, impl_->balance != 0, st);
                else launch_wstep<MEVO_WCAPS_A, 16, 1>(prm, impl_->n_sm, impl_->balance != 0, st);
                break;
            case 1: launch_wstep<MEVO_WCAPS_C, 12, 1>(prm, impl_->n_sm, impl_->balance != 0, st); break;
            case 2: launch_wstep<MEVO_WCAPS_B, 10, 1>(prm, impl_->n_sm, impl_->balance != 0, st); break;
            default: gstep_dispatch(c, impl_->class_ppt[c] | impl_->class_xd[c], GAction::Launch, &prm, impl_->class_smem[c], st); break;
        }
        CUDA_CHECK(cudaGetLastError());
        stats_.launches++;
    }
    for (auto& e : envs_) e.host_fresh = false;
}

int Batch::step(int n_steps, bool apply_actions, uint8_t* unstable_host, int n_substeps) {
    step_async(n_steps, apply_actions, n_substeps);
    const int n = num_envs();
    CUDA_CHECK(cudaMemcpyAsync(impl_->unstable_host, d_unstable_, n, cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    int count = 0;
    for (int i = 0; i < n; i++) count += impl_->unstable_host[i] ? 1 : 0;
    if (unstable_host) std::memcpy(unstable_host, impl_->unstable_host, n);
    return count;
}

// ------------------------------------------------------------------------------------------------
// CUDA graphs

void Batch::capture_begin() {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (capturing_) throw std::runtime_error("capture_begin: a capture is already open");
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamBeginCapture((cudaStream_t)stream_, cudaStreamCaptureModeThreadLocal));
    capturing_ = true;
}

int Batch::capture_end() {
    if (!capturing_) throw std::runtime_error("capture_end without capture_begin");
    capturing_ = false;
    cudaGraph_t graph = nullptr;
    CUDA_CHECK(cudaStreamEndCapture((cudaStream_t)stream_, &graph));
    cudaGraphExec_t exec = nullptr;
    const cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    CUDA_CHECK(e);
    graphs_.push_back(exec);
    return (int)graphs_.size() - 1;
}

void Batch::graph_launch(int id) {
    if (id < 0 || id >= (int)graphs_.size() || !graphs_[id]) throw std::runtime_error("graph_launch: no such graph");
    CUDA_CHECK(cudaGraphLaunch((cudaGraphExec_t)graphs_[id], (cudaStream_t)stream_));
    for (auto& e : envs_) e.host_fresh = false;
}

void Batch::graph_destroy(int id) {
    if (id < 0 || id >= (int)graphs_.size() || !graphs_[id]) return;
    cudaGraphExecDestroy((cudaGraphExec_t)graphs_[id]);
    graphs_[id] = nullptr;
}

// ------------------------------------------------------------------------------------------------
// reset snapshot

void Batch::save_reset_state() {
    download_all();
    for (int i = 0; i < num_envs(); i++) { impl_->pos0[i] = envs_[i].pos; impl_->vel0[i] = envs_[i].vel; }
    CUDA_CHECK(cudaSetDevice(device_));
    // pos/vel sections of the device snapshot = current state
    CUDA_CHECK(cudaMemcpyAsync(d_state0_, d_state_, sizeof(double) * state_doubles_, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

void Batch::reset_envs(const int* env_ids, int n) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (n <= 0) return;
    for (int i = 0; i < n; i++)
        if (env_ids[i] < 0 || env_ids[i] >= num_envs()) throw std::runtime_error("reset: environment id out of range");
    if (n > impl_->ids_cap) {                       // id list on the device, grown on demand and kept
        cudaFree(impl_->d_ids); impl_->d_ids = nullptr; impl_->ids_cap = 0;
        CUDA_CHECK(cudaMalloc(&impl_->d_ids, sizeof(int) * n));
        impl_->ids_cap = n;
    }
    CUDA_CHECK(cudaMemcpyAsync(impl_->d_ids, env_ids, sizeof(int) * n, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    reset_kernel<<<n, 64, 0, (cudaStream_t)stream_>>>(d_state_, d_state0_, reinterpret_cast<const EnvDesc*>(d_envdesc_), d_design_, impl_->d_ids, n,
                                                       d_time_, d_unstable_);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));       // env_ids is the caller's (pageable) memory
    for (int i = 0; i < n; i++) envs_[env_ids[i]].host_fresh = false;
}

void Batch::reset_all() {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    reset_kernel<<<num_envs(), 64, 0, (cudaStream_t)stream_>>>(d_state_, d_state0_, reinterpret_cast<const EnvDesc*>(d_envdesc_), d_design_,
                                                                nullptr, num_envs(), d_time_, d_unstable_);
    CUDA_CHECK(cudaGetLastError());
    for (auto& e : envs_) e.host_fresh = false;
}

// ---- batched snapshot history (Environment::save_snapshot / revert_to_snapshot / Sim::revert, Environment.cpp:239-263,
// Sim.cpp:126-133): slot 0 is the reset snapshot (save_reset_state / reset_*), slots 1 .. kSlots-1 are the caller's.

void Batch::save_snapshot(int slot) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (slot == 0) { save_reset_state(); return; }
    if (slot < 0 || slot >= Impl::kSlots) throw std::runtime_error("save_snapshot: slot out of range (0 .. 7)");
    if (!impl_->d_slot[slot]) {
        CUDA_CHECK(cudaMalloc(&impl_->d_slot[slot], sizeof(double) * (state_doubles_ + 2)));
        CUDA_CHECK(cudaMalloc(&impl_->d_slot_time[slot], sizeof(int) * num_envs()));
    }
    cudaStream_t st = (cudaStream_t)stream_;
    CUDA_CHECK(cudaMemcpyAsync(impl_->d_slot[slot], d_state_, sizeof(double) * state_doubles_, cudaMemcpyDeviceToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(impl_->d_slot_time[slot], d_time_, sizeof(int) * num_envs(), cudaMemcpyDeviceToDevice, st));
}

void Batch::revert_envs(const int* env_ids, int n, int slot) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (slot == 0) {
        if (env_ids) reset_envs(env_ids, n); else reset_all();
        return;
    }
    if (slot < 0 || slot >= Impl::kSlots) throw std::runtime_error("revert: slot out of range (0 .. 7)");
    if (!impl_->d_slot[slot]) {
        // the reference prints and does nothing when no snapshot exists for the time asked (Environment.cpp:249-252)
        std::printf("Error: Could not revert simulation - no save found for time argument.\n");
        return;
    }
    cudaStream_t st = (cudaStream_t)stream_;
    const EnvDesc* ed = reinterpret_cast<const EnvDesc*>(d_envdesc_);
    if (!env_ids) {
        revert_kernel<<<num_envs(), 64, 0, st>>>(d_state_, impl_->d_slot[slot], impl_->d_slot_time[slot], ed, d_design_, nullptr, num_envs(),
                                                 d_time_, d_unstable_);
        CUDA_CHECK(cudaGetLastError());
        for (auto& e : envs_) e.host_fresh = false;
        return;
    }
    if (n <= 0) return;
    for (int i = 0; i < n; i++)
        if (env_ids[i] < 0 || env_ids[i] >= num_envs()) throw std::runtime_error("revert: environment id out of range");
    if (n > impl_->ids_cap) {
        cudaFree(impl_->d_ids); impl_->d_ids = nullptr; impl_->ids_cap = 0;
        CUDA_CHECK(cudaMalloc(&impl_->d_ids, sizeof(int) * n));
        impl_->ids_cap = n;
    }
    CUDA_CHECK(cudaMemcpyAsync(impl_->d_ids, env_ids, sizeof(int) * n, cudaMemcpyHostToDevice, st));
    revert_kernel<<<n, 64, 0, st>>>(d_state_, impl_->d_slot[slot], impl_->d_slot_time[slot], ed, d_design_, impl_->d_ids, n, d_time_, d_unstable_);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));          // env_ids is the caller's (pageable) memory
    for (int i = 0; i < n; i++) envs_[env_ids[i]].host_fresh = false;
}

// Sim::pos_at_time / vel_at_time of one environment from a slot (global numbering; points of pruned all-FIXED objects never move
// and read their construction positions, velocity 0).  Returns the slot's time for that environment, or -1 for an empty slot.
long Batch::read_snapshot(int env, int slot, double* pos, double* vel) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (slot < 0 || slot >= Impl::kSlots) throw std::runtime_error("read_snapshot: slot out of range (0 .. 7)");
    const double* src = slot == 0 ? d_state0_ : impl_->d_slot[slot];
    if (!src) return -1;
    download_state(env);                            // host mirror: static points
    const EnvHost& e = envs_.at(env);
    const int Pd = (int)e.dyn_points.size();
    std::vector<double> s(4 * (size_t)Pd);
    int t = 0;
    cudaStream_t st = (cudaStream_t)stream_;
    if (Pd) CUDA_CHECK(cudaMemcpyAsync(s.data(), src + e.state_off, sizeof(double) * s.size(), cudaMemcpyDeviceToHost, st));
    if (slot != 0) CUDA_CHECK(cudaMemcpyAsync(&t, impl_->d_slot_time[slot] + env, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (pos) std::memcpy(pos, e.world.pos.data(), sizeof(double) * e.pos.size());
    if (vel) std::fill(vel, vel + e.pos.size(), 0.0);
    for (int k = 0; k < Pd; k++) {
        const int g = e.dyn_points[k];
        for (int c = 0; c < 2; c++) {
            if (pos) pos[2 * g + c] = s[2 * k + c];
            if (vel) vel[2 * g + c] = s[2 * Pd + 2 * k + c];
        }
    }
    return t;
}

void Batch::object_com(int obj_index, double* out_host_2n) {
    object_com_device(obj_index, nullptr);
    CUDA_CHECK(cudaMemcpyAsync(out_host_2n, impl_->d_com_out, sizeof(double) * 2 * num_envs(), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

// One host round trip for what a rollout loop reads after a step: the unstable flags of the last step launch and the centre of
// mass of object `obj_index` in every environment (either output may be null).  Enqueues the COM kernel and both copies on the
// batch's stream and waits once.
int Batch::fetch_step_results(uint8_t* unstable_host, int obj_index, double* com_host_2n) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    const int n = num_envs();
    cudaStream_t st = (cudaStream_t)stream_;
    if (com_host_2n) {
        object_com_device(obj_index, nullptr);
        CUDA_CHECK(cudaMemcpyAsync(com_host_2n, impl_->d_com_out, sizeof(double) * 2 * n, cudaMemcpyDeviceToHost, st));
    }
    CUDA_CHECK(cudaMemcpyAsync(impl_->unstable_host, d_unstable_, n, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    int count = 0;
    for (int i = 0; i < n; i++) count += impl_->unstable_host[i] ? 1 : 0;
    if (unstable_host) std::memcpy(unstable_host, impl_->unstable_host, n);
    return count;
}

void Batch::bind_object_ranges(int obj_index) {
    const int n = num_envs();
    if (impl_->com_obj == obj_index) return;
    std::vector<int> ranges(3 * (size_t)n, 0);
    for (int i = 0; i < n; i++) {
        const EnvHost& e = envs_[i];
        ranges[3 * i + 2] = (int)e.dyn_points.size();
        if (obj_index >= (int)e.world.objects.size()) continue;
        const VoxelObject& ob = e.world.objects[obj_index];
        const int l0 = e.point_local[ob.min_point];
        if (l0 < 0) continue;   // static object: not tracked on the device
        ranges[3 * i] = l0;
        ranges[3 * i + 1] = ob.max_point - ob.min_point + 1;
    }
    cudaFree(impl_->d_com_ranges); cudaFree(impl_->d_com_out);
    CUDA_CHECK(cudaMalloc(&impl_->d_com_ranges, sizeof(int) * ranges.size()));
    CUDA_CHECK(cudaMalloc(&impl_->d_com_out, sizeof(double) * 2 * n));
    CUDA_CHECK(cudaMemcpy(impl_->d_com_ranges, ranges.data(), sizeof(int) * ranges.size(), cudaMemcpyHostToDevice));
    impl_->com_obj = obj_index;
}

void Batch::object_com_device(int obj_index, double* out_device_2n) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    const int n = num_envs();
    bind_object_ranges(obj_index);
    com_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(d_state_, reinterpret_cast<const EnvDesc*>(d_envdesc_), impl_->d_com_ranges, n,
                                                                   out_device_2n ? out_device_2n : impl_->d_com_out);
    CUDA_CHECK(cudaGetLastError());
}

void Batch::object_orientation_strided_device(int obj_index, void* out, int row_stride, bool is_f32, const double* state_like) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    const int n = num_envs();
    bind_object_ranges(obj_index);
    const EnvDesc* ed = reinterpret_cast<const EnvDesc*>(d_envdesc_);
    if (is_f32)
        orientation_kernel<float><<<(n + 3) / 4, 128, 0, (cudaStream_t)stream_>>>(state_like ? state_like : d_state_, d_state0_, ed, impl_->d_com_ranges, n,
                                                                                 static_cast<float*>(out), row_stride);
    else
        orientation_kernel<double><<<(n + 3) / 4, 128, 0, (cudaStream_t)stream_>>>(state_like ? state_like : d_state_, d_state0_, ed, impl_->d_com_ranges, n,
                                                                                  static_cast<double*>(out), row_stride);
    CUDA_CHECK(cudaGetLastError());
}

void Batch::object_orientation_device(int obj_index, double* out_device_n) {
    bind_object_ranges(obj_index);                 // allocates d_com_out
    object_orientation_strided_device(obj_index, out_device_n ? out_device_n : impl_->d_com_out, 1, false);
}

void Batch::object_orientation(int obj_index, double* out_host_n) {
    object_orientation_device(obj_index, nullptr);
    CUDA_CHECK(cudaMemcpyAsync(out_host_n, impl_->d_com_out, sizeof(double) * num_envs(), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

int Batch::object_obs_width(int obj_index) const {
    int w = 2;
    for (const EnvHost& e : envs_) {
        if (obj_index >= (int)e.world.objects.size()) continue;
        const VoxelObject& ob = e.world.objects[obj_index];
        w = std::max(w, 2 + 2 * (ob.max_point - ob.min_point + 1));
    }
    return w;
}

void Batch::object_obs_device(int obj_index, double* out_device, int stride) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    const int n = num_envs();
    bind_object_ranges(obj_index);
    obs_kernel<<<(n + 3) / 4, 128, 0, (cudaStream_t)stream_>>>(d_state_, reinterpret_cast<const EnvDesc*>(d_envdesc_), impl_->d_com_ranges, n,
                                                              out_device, stride);
    CUDA_CHECK(cudaGetLastError());
}

void Batch::object_obs(int obj_index, double* out_host, int stride) {
    const size_t bytes = sizeof(double) * (size_t)stride * num_envs();
    double* d = nullptr;
    CUDA_CHECK(cudaSetDevice(device_));
    CUDA_CHECK(cudaMalloc(&d, bytes));
    try {
        object_obs_device(obj_index, d, stride);
        CUDA_CHECK(cudaMemcpyAsync(out_host, d, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    } catch (...) { cudaFree(d); throw; }
    cudaFree(d);
}

// ------------------------------------------------------------------------------------------------
// debug hooks

void Batch::enable_candidate_dump(int cap_per_env) {
    CUDA_CHECK(cudaSetDevice(device_));
    cudaFree(d_cand_); cudaFree(d_cand_n_);
    d_cand_ = nullptr; d_cand_n_ = nullptr;
    cand_cap_ = cap_per_env;
    if (cap_per_env <= 0 || num_envs() == 0) return;
    CUDA_CHECK(cudaMalloc(&d_cand_, sizeof(unsigned int) * (size_t)cap_per_env * num_envs()));
    CUDA_CHECK(cudaMalloc(&d_cand_n_, sizeof(int) * num_envs()));
    CUDA_CHECK(cudaMemset(d_cand_n_, 0, sizeof(int) * num_envs()));
}

void Batch::enable_force_dump() {
    CUDA_CHECK(cudaSetDevice(device_));
    force_dump_ = true;
    if (!d_state_) return;                          // not finalized yet: finalize() allocates it
    cudaFree(impl_->d_cf); impl_->d_cf = nullptr;
    CUDA_CHECK(cudaMalloc(&impl_->d_cf, sizeof(double) * (state_doubles_ + 2)));
    CUDA_CHECK(cudaMemset(impl_->d_cf, 0, sizeof(double) * (state_doubles_ + 2)));
}

void Batch::read_collision_forces(int env, double* out) {
    // collision_forces (PhysicsEngine.cpp:715-751) after the collision phase of the first substep of the last step call, global
    // numbering; points of pruned all-FIXED objects read 0 (the engine never accumulates forces on points that cannot move)
    if (!impl_->d_cf) throw std::runtime_error("read_collision_forces: enable_force_dump() first");
    CUDA_CHECK(cudaSetDevice(device_));
    sync();
    const EnvHost& e = envs_.at(env);
    const int Pd = (int)e.dyn_points.size();
    std::vector<double> cfv(2 * (size_t)Pd);
    if (Pd) CUDA_CHECK(cudaMemcpy(cfv.data(), impl_->d_cf + e.state_off, sizeof(double) * cfv.size(), cudaMemcpyDeviceToHost));
    std::fill(out, out + e.pos.size(), 0.0);
    for (int k = 0; k < Pd; k++) {
        out[2 * e.dyn_points[k]] = cfv[2 * k];
        out[2 * e.dyn_points[k] + 1] = cfv[2 * k + 1];
    }
}

void Batch::leaf_info(int env, int dynamic, int leaf, int* obj, int* boxel) const {
    // leaves are numbered per environment: dynamic leaves over dynamic objects in load order, static leaves
    // over static objects in load order; inside an object by ascending boxel index
    const EnvHost& e = envs_.at(env);
    int k = 0;
    for (int o = 0; o < (int)e.world.objects.size(); o++) {
        const VoxelObject& ob = e.world.objects[o];
        const bool is_dyn = !(ob.all_fixed && !ob.is_robot) || e.obj_force_dynamic[o];
        if ((is_dyn ? 1 : 0) != (dynamic ? 1 : 0)) continue;
        if (leaf < k + (int)ob.surface_boxels.size()) { *obj = o; *boxel = ob.surface_boxels[leaf - k]; return; }
        k += (int)ob.surface_boxels.size();
    }
    *obj = -1; *boxel = -1;
}

int Batch::read_candidates(int env, int* out_pairs, int cap) {
    if (!d_cand_) return -1;
    CUDA_CHECK(cudaSetDevice(device_));
    sync();
    int n = 0;
    CUDA_CHECK(cudaMemcpy(&n, d_cand_n_ + env, sizeof(int), cudaMemcpyDeviceToHost));
    const int m = std::min(n, cand_cap_);
    std::vector<unsigned int> codes(m);
    if (m) CUDA_CHECK(cudaMemcpy(codes.data(), d_cand_ + (size_t)env * cand_cap_, sizeof(unsigned int) * m, cudaMemcpyDeviceToHost));
    int w = 0;
    for (int i = 0; i < m && w < cap; i++, w++) {
        const unsigned c = codes[i];
        int oa, ba, ob, bb;
        leaf_info(env, 1, (int)(c & kCodeLeafMask), &oa, &ba);
        leaf_info(env, (c & kCodeStatic) ? 0 : 1, (int)((c >> 12) & kCodeLeafMask), &ob, &bb);
        out_pairs[4 * w] = oa; out_pairs[4 * w + 1] = ba; out_pairs[4 * w + 2] = ob; out_pairs[4 * w + 3] = bb;
    }
    // pairs of leaves that share a point are candidates in every substep by construction (their boxes both
    // contain the shared point); the kernel never tests them, so they are appended here
    const std::vector<int>& ap = envs_.at(env).always_pairs;
    for (size_t i = 0; i + 1 < ap.size() && w < cap; i += 2, w++) {
        int oa, ba, ob, bb;
        leaf_info(env, 1, ap[i], &oa, &ba);
        leaf_info(env, 1, ap[i + 1], &ob, &bb);
        out_pairs[4 * w] = oa; out_pairs[4 * w + 1] = ba; out_pairs[4 * w + 2] = ob; out_pairs[4 * w + 3] = bb;
    }
    return n + (int)ap.size() / 2;
}

int Batch::error_flags(int env) {
    CUDA_CHECK(cudaSetDevice(device_));
    sync();
    int v = 0;
    CUDA_CHECK(cudaMemcpy(&v, d_err_ + env, sizeof(int), cudaMemcpyDeviceToHost));
    return v;
}

}  // namespace mevo
