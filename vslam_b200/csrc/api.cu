// Host side of the C ABI (include/vslam_b200.h): context, device frame cache, batched alignment driver.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "solve.cuh"

namespace
{
thread_local std::string g_err;
int fail(int code, const std::string & msg)
{
  g_err = msg;
  return code;
}
#define CK(call)                                                                                          \
  do {                                                                                                    \
    cudaError_t e_ = (call);                                                                              \
    if (e_ != cudaSuccess)                                                                                \
      return fail(VSLAMB200_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                   \
  } while (0)

struct DevBuf
{
  void * p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes, bool zero = false)
  {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) return e;
    cap = want;
    if (zero) e = cudaMemset(p, 0, want);
    return e;
  }
  void release()
  {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T * as()
  {
    return reinterpret_cast<T *>(p);
  }
};

struct Arena
{
  void * ptr = nullptr;
  size_t bytes = 0;
  int refs = 0;
  bool spare = false;  // no live frames; kept for the next upload of the same size (no cudaMalloc/cudaFree per batch)
  // asynchronous upload: `ready` is recorded on the copy stream behind the copies and the preprocessing kernels;
  // work that reads the frames waits for it on the compute stream
  cudaEvent_t ready = nullptr;
  bool pending = false;
};

struct FrameHost
{
  bool used = false;
  uint64_t id = 0;
  int arena = -1;
  vb::FrameDev dev;
};

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr int kMinJobsPerPart = 96;  // below this a half batch no longer fills the GPU
}  // namespace

namespace vb
{
int set_last_error(int code, const char * msg) { return fail(code, msg); }
}  // namespace vb

struct vslamb200_ctx
{
  int device = 0;
  cudaStream_t own = nullptr, stream = nullptr;
  vslamb200_config cfg;
  std::unordered_map<uint64_t, int> idToSlot;
  std::vector<FrameHost> frames;
  std::vector<int> freeSlots;
  std::vector<Arena> arenas;
  DevBuf dFrames;  // FrameDev table
  DevBuf stageGray, stageDepth, stageU16, slotsBuf;
  // alignment workspace
  DevBuf jobs, views, chunkCounts, partials, lists, trace, plane, ext;
  int traceEnabled = 0, timing = 0;
  int traceJobs = 0, traceLevels = 0, traceMaxIter = 0;
  vslamb200_stats stats;
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  // asynchronous uploads: copies run on their own stream; the slot lists of the preprocessing launches live in a
  // small ring of pinned buffers so that no launch needs a host synchronisation
  cudaStream_t copyStream = nullptr;
  cudaEvent_t copyDone = nullptr, computeMark = nullptr;
  // asynchronous uploads: copies on copyStream; conversion, pyramids and gradients on prepStream behind copyDone, so the
  // next transfer does not queue behind them; u16Done: the staged 16-bit depth of the last upload has been consumed
  cudaStream_t prepStream = nullptr;
  cudaEvent_t u16Done = nullptr;
  bool u16Pending = false;
  static constexpr int kSlotRing = 4;
  int * hSlots[kSlotRing] = {nullptr, nullptr, nullptr, nullptr};
  size_t hSlotsCap[kSlotRing] = {0, 0, 0, 0};
  DevBuf dSlots[kSlotRing];
  // pinned staging for the job / view tables of an alignment call (read by a copy kernel, see launch_stage_copy)
  void * hStage = nullptr;
  size_t hStageCap = 0;
  // large batches run as two halves on two streams (see vslamb200_align)
  static constexpr int kMaxParts = 8;
  cudaStream_t part[kMaxParts] = {};
  cudaEvent_t forkEv = nullptr, joinEv[kMaxParts] = {};
  cudaEvent_t slotsEv[kSlotRing] = {nullptr, nullptr, nullptr, nullptr};
  unsigned slotsNext = 0;
};

namespace
{
void level_dims(int W, int H, int nLevels, vb::FrameDev & f)
{
  f.nLevels = nLevels;
  f.W[0] = f.Wd[0] = W;
  f.H[0] = f.Hd[0] = H;
  for (int l = 1; l < nLevels; l++) {
    f.W[l] = (f.W[l - 1] + 1) / 2;  // cv::pyrDown (Frame.cpp:264)
    f.H[l] = (f.H[l - 1] + 1) / 2;
    f.Wd[l] = (int)(f.Wd[l - 1] * 0.5);  // algorithm::resize (algorithm.h:93-94)
    f.Hd[l] = (int)(f.Hd[l - 1] * 0.5);
  }
}

// per-frame block layout; returns bytes.  base == nullptr only sizes.
size_t layout_frame(vb::FrameDev & f, uint8_t * base)
{
  size_t off = 0;
  auto take = [&](size_t bytes) {
    uint8_t * p = base ? base + off : nullptr;
    off += align_up(std::max<size_t>(bytes, 1), 256);
    return p;
  };
  for (int l = 0; l < f.nLevels; l++) {
    const size_t n = (size_t)f.W[l] * f.H[l], nd = (size_t)f.Wd[l] * f.Hd[l];
    f.I[l] = take(n);
    f.Z[l] = (float *)take(nd * 4);
    f.Zsf[l] = l == 0 ? f.Z[0] : (float *)take(nd * 4);
    f.gx[l] = (int16_t *)take(n * 2);
    f.gy[l] = (int16_t *)take(n * 2);
  }
  return off;
}

int release_slot(vslamb200_ctx * c, int slot)
{
  FrameHost & f = c->frames[slot];
  if (!f.used) return 0;
  c->idToSlot.erase(f.id);
  f.used = false;
  Arena & a = c->arenas[f.arena];
  if (--a.refs == 0) a.spare = true;
  c->freeSlots.push_back(slot);
  return 0;
}

// frees spare arenas, keeping the `keep` largest
void trim_spares(vslamb200_ctx * c, int keep)
{
  std::vector<int> sp;
  for (size_t i = 0; i < c->arenas.size(); i++)
    if (c->arenas[i].ptr && c->arenas[i].spare) sp.push_back((int)i);
  std::sort(sp.begin(), sp.end(), [&](int x, int y) { return c->arenas[x].bytes > c->arenas[y].bytes; });
  for (size_t k = keep; k < sp.size(); k++) {
    Arena & a = c->arenas[sp[k]];
    if (a.pending) cudaEventSynchronize(a.ready);
    if (a.ready) cudaEventDestroy(a.ready);
    cudaFree(a.ptr);
    a = Arena();
  }
}

int upload_table(vslamb200_ctx * c, const std::vector<int> & slots)
{
  const size_t need = c->frames.size() * sizeof(vb::FrameDev);
  if (need > c->dFrames.cap) {
    // grow: re-upload every live entry
    CK(cudaStreamSynchronize(c->stream));
    CK(c->dFrames.ensure(std::max<size_t>(need * 2, 64 * sizeof(vb::FrameDev))));
    std::vector<vb::FrameDev> all(c->frames.size());
    for (size_t i = 0; i < c->frames.size(); i++) all[i] = c->frames[i].dev;
    CK(cudaMemcpyAsync(c->dFrames.p, all.data(), need, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
  }
  // contiguous runs
  size_t i = 0;
  std::vector<int> s(slots);
  std::sort(s.begin(), s.end());
  std::vector<vb::FrameDev> tmp;
  while (i < s.size()) {
    size_t j = i;
    while (j + 1 < s.size() && s[j + 1] == s[j] + 1) j++;
    tmp.resize(j - i + 1);
    for (size_t k = i; k <= j; k++) tmp[k - i] = c->frames[s[k]].dev;
    CK(cudaMemcpyAsync(c->dFrames.as<vb::FrameDev>() + s[i], tmp.data(), tmp.size() * sizeof(vb::FrameDev),
                       cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // tmp is reused
    i = j + 1;
  }
  return 0;
}

// allocate slots + one arena for n frames of identical geometry
// borrowGray / borrowDepth (both or neither): the level-0 planes are the caller's tightly packed device buffers, used in
// place (vslamb200_frames_attach_device)
int make_frames(vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const double * cams, int nLevels,
                std::vector<int> & slots, const uint8_t * borrowGray = nullptr, const float * borrowDepth = nullptr)
{
  if (n <= 0 || !ids || !cams) return fail(VSLAMB200_EINVAL, "frames_upload: bad arguments");
  if (nLevels < 1 || nLevels > vb::MAX_LEVELS) return fail(VSLAMB200_EINVAL, "frames_upload: n_levels out of range");
  if (W < 1 || H < 1 || W > 4095 || H > 4095) return fail(VSLAMB200_EINVAL, "frames_upload: image size out of range");
  for (int i = 0; i < n; i++)
    if (std::abs(W / 2 - cams[4 * i + 2]) > 10)  // Frame.cpp:150-155
      return fail(VSLAMB200_EDIMS, "Inconsistent camera parameters / image / depth dimensions detected.");
  vb::FrameDev proto;
  std::memset(&proto, 0, sizeof(proto));
  level_dims(W, H, nLevels, proto);
  const size_t frameBytes = layout_frame(proto, nullptr);
  {
    std::vector<uint64_t> sorted(ids, ids + n);
    std::sort(sorted.begin(), sorted.end());
    if (std::adjacent_find(sorted.begin(), sorted.end()) != sorted.end())
      return fail(VSLAMB200_EINVAL, "frames_upload: the same frame id appears twice in one call");
  }
  // Frames re-uploaded under the same id are replaced.  Work of the compute stream that still uses the old frame is
  // ordered before the new upload on that stream; an asynchronous upload of the old frame that still runs on the copy
  // stream (its preprocessing reads the frame table entry this call is about to overwrite) is waited for first.
  for (int i = 0; i < n; i++) {
    auto it = c->idToSlot.find(ids[i]);
    if (it == c->idToSlot.end()) continue;
    Arena & old = c->arenas[c->frames[it->second].arena];
    if (old.pending) {
      CK(cudaStreamWaitEvent(c->stream, old.ready, 0));  // later copy-stream work is behind it by stream order
    }
    release_slot(c, it->second);
  }
  const size_t need = frameBytes * n;
  int arenaIdx = -1;
  for (size_t i = 0; i < c->arenas.size(); i++) {  // smallest spare arena that fits without wasting half of it
    const Arena & s = c->arenas[i];
    if (s.ptr && s.spare && s.bytes >= need && s.bytes <= 2 * need &&
        (arenaIdx < 0 || s.bytes < c->arenas[arenaIdx].bytes))
      arenaIdx = (int)i;
  }
  if (arenaIdx < 0) {
    Arena a;
    a.bytes = need;
    if (cudaMalloc(&a.ptr, a.bytes) != cudaSuccess) {
      cudaGetLastError();
      CK(cudaStreamSynchronize(c->stream));
      trim_spares(c, 0);
      CK(cudaMalloc(&a.ptr, a.bytes));
    }
    for (size_t i = 0; i < c->arenas.size(); i++)
      if (!c->arenas[i].ptr) arenaIdx = (int)i;
    if (arenaIdx < 0) {
      c->arenas.push_back(a);
      arenaIdx = (int)c->arenas.size() - 1;
    } else {
      c->arenas[arenaIdx] = a;
    }
  }
  Arena & a = c->arenas[arenaIdx];
  if (a.pending) CK(cudaStreamWaitEvent(c->stream, a.ready, 0));  // a recycled arena whose asynchronous upload still runs
  a.spare = false;
  a.refs = n;
  trim_spares(c, 2);
  slots.resize(n);
  for (int i = 0; i < n; i++) {
    int slot;
    if (!c->freeSlots.empty()) {
      slot = c->freeSlots.back();
      c->freeSlots.pop_back();
    } else {
      slot = (int)c->frames.size();
      c->frames.emplace_back();
    }
    FrameHost & f = c->frames[slot];
    f.used = true;
    f.id = ids[i];
    f.arena = arenaIdx;
    f.dev = proto;
    layout_frame(f.dev, (uint8_t *)a.ptr + frameBytes * i);
    if (borrowGray) {
      f.dev.I[0] = const_cast<uint8_t *>(borrowGray) + (size_t)W * H * i;
      f.dev.Z[0] = f.dev.Zsf[0] = const_cast<float *>(borrowDepth) + (size_t)W * H * i;
    }
    for (int l = 0; l < nLevels; l++) {
      const double s = std::pow(0.5, l);  // Camera::resize (Camera.cpp:53-57, Frame.cpp:267)
      for (int k = 0; k < 4; k++) f.dev.cam[l][k] = cams[4 * i + k] * s;
    }
    c->idToSlot[ids[i]] = slot;
    slots[i] = slot;
  }
  return upload_table(c, slots);
}

// the compute stream waits for the asynchronous upload of the frame in `slot`, if one is still pending
int wait_frame_ready(vslamb200_ctx * c, int slot)
{
  Arena & a = c->arenas[c->frames[slot].arena];
  if (!a.pending) return 0;
  if (cudaEventQuery(a.ready) == cudaSuccess) {
    a.pending = false;
    return 0;
  }
  cudaGetLastError();  // cudaErrorNotReady is not sticky, but clear it
  CK(cudaStreamWaitEvent(c->stream, a.ready, 0));
  return 0;
}

// pyramid + gradients for uploaded level-0 planes; queued on `stream`, host-synchronised only if asked
// stagedU16 != nullptr: the level-0 depth planes are first filled from 16-bit sensor depth staged on the device
int preprocess(vslamb200_ctx * c, const std::vector<int> & slots, int nLevels, bool sync, cudaStream_t stream,
               const uint16_t * stagedU16 = nullptr, double divisor = 1.0, cudaEvent_t afterU16 = nullptr)
{
  const int n = (int)slots.size();
  const unsigned ring = c->slotsNext++ % vslamb200_ctx::kSlotRing;
  CK(cudaEventSynchronize(c->slotsEv[ring]));  // the launches that last read this ring entry are done
  if ((size_t)n > c->hSlotsCap[ring]) {
    if (c->hSlots[ring]) cudaFreeHost(c->hSlots[ring]);
    c->hSlots[ring] = nullptr;
    c->hSlotsCap[ring] = 0;
    CK(cudaHostAlloc((void **)&c->hSlots[ring], sizeof(int) * (size_t)n * 2, cudaHostAllocDefault));
    c->hSlotsCap[ring] = (size_t)n * 2;
  }
  CK(c->dSlots[ring].ensure(n * sizeof(int)));
  std::memcpy(c->hSlots[ring], slots.data(), n * sizeof(int));
  CK(cudaMemcpyAsync(c->dSlots[ring].p, c->hSlots[ring], n * sizeof(int), cudaMemcpyHostToDevice, stream));
  const int * dSlots = c->dSlots[ring].as<int>();
  const vb::FrameDev & f = c->frames[slots[0]].dev;
  const vb::FrameDev * table = c->dFrames.as<vb::FrameDev>();
  if (c->timing) CK(cudaEventRecord(c->ev[0], stream));
  if (stagedU16) {
    vb::launch_depth_u16(stagedU16, divisor, f.W[0], f.H[0], table, dSlots, n, stream);
    c->stats.kernel_launches += 1;
    if (afterU16) CK(cudaEventRecord(afterU16, stream));  // the staging buffer may be overwritten from here on
  }
  for (int l = 1; l < nLevels; l++) {
    vb::launch_pyramid_level(table, dSlots, n, l, f.W[l], f.H[l], f.Wd[l], f.Hd[l], f.W[l - 1], f.H[l - 1], f.Wd[l - 1],
                             f.Hd[l - 1], stream);
    c->stats.kernel_launches += 2;
  }
  for (int l = 0; l < nLevels; l++) {
    vb::launch_gradients(table, dSlots, n, l, f.W[l], f.H[l], stream);
    c->stats.kernel_launches += 1;
  }
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->slotsEv[ring], stream));
  if (c->timing) {
    CK(cudaEventRecord(c->ev[1], stream));
    CK(cudaStreamSynchronize(stream));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->stats.ms_preprocess += ms;
  }
  if (sync) CK(cudaStreamSynchronize(stream));
  c->stats.frames_uploaded += n;
  return 0;
}
}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI: context and frames
// ------------------------------------------------------------------------------------------------
extern "C" {

const char * vslamb200_last_error(void) { return g_err.c_str(); }
int vslamb200_version(void) { return 100; }

static int check_cfg(const vslamb200_config * cfg)
{
  if (!cfg) return fail(VSLAMB200_EINVAL, "config is null");
  if (cfg->loss_kind < 0 || cfg->loss_kind > VSLAMB200_LOSS_TDIST) return fail(VSLAMB200_EINVAL, "bad loss_kind");
  if (cfg->scaler_kind < 0 || cfg->scaler_kind > VSLAMB200_SCALER_TDIST) return fail(VSLAMB200_EINVAL, "bad scaler_kind");
  if (cfg->max_iterations < 1 || cfg->max_iterations > 100000) return fail(VSLAMB200_EINVAL, "bad max_iterations");
  if (cfg->solver_kind != VSLAMB200_SOLVER_GN && cfg->solver_kind != VSLAMB200_SOLVER_LM)
    return fail(VSLAMB200_EINVAL, "bad solver_kind");
  if (cfg->solver_kind == VSLAMB200_SOLVER_LM &&
      !(cfg->lm_lambda0 > 0.0 && cfg->lm_up > 1.0 && cfg->lm_down > 0.0 && cfg->lm_down <= 1.0))
    return fail(VSLAMB200_EINVAL, "bad Levenberg-Marquardt parameters (lambda0 > 0, up > 1, 0 < down <= 1)");
  if (cfg->sampling != VSLAMB200_SAMPLING_NEAREST && cfg->sampling != VSLAMB200_SAMPLING_BILINEAR)
    return fail(VSLAMB200_EINVAL, "bad sampling");
  if (!(cfg->depth_weight >= 0.0)) return fail(VSLAMB200_EINVAL, "bad depth_weight");
  if (cfg->chi2_mode != VSLAMB200_CHI2_REFERENCE && cfg->chi2_mode != VSLAMB200_CHI2_WEIGHTED)
    return fail(VSLAMB200_EINVAL, "bad chi2_mode");
  if (cfg->level_kernel < 0 || cfg->level_kernel > VSLAMB200_KERNEL_MOMENTS) return fail(VSLAMB200_EINVAL, "bad level_kernel");
  if (cfg->cluster_size != 0 && cfg->cluster_size != 1 && cfg->cluster_size != 2 && cfg->cluster_size != 4 &&
      cfg->cluster_size != 8 && cfg->cluster_size != 16)
    return fail(VSLAMB200_EINVAL, "bad cluster_size (0, 1, 2, 4, 8 or 16)");
  if (cfg->stream_parts < 0 || cfg->stream_parts > 8) return fail(VSLAMB200_EINVAL, "bad stream_parts (0..8)");
  if (cfg->sampling == VSLAMB200_SAMPLING_NEAREST && cfg->depth_weight != 0.0)
    return fail(VSLAMB200_EINVAL, "the depth residual needs bilinear sampling");
  if (cfg->sampling == VSLAMB200_SAMPLING_BILINEAR && cfg->loss_kind != VSLAMB200_LOSS_NONE &&
      cfg->scaler_kind != VSLAMB200_SCALER_IDENTITY && cfg->scaler_kind != VSLAMB200_SCALER_TDIST)
    return fail(VSLAMB200_EINVAL, "bilinear sampling supports the identity and t-distribution scalers");
  return 0;
}

int vslamb200_create(const vslamb200_config * cfg, int device, vslamb200_ctx ** out)
{
  if (!out) return fail(VSLAMB200_EINVAL, "out is null");
  if (int e = check_cfg(cfg)) return e;
  int nDev = 0;
  cudaError_t ce = cudaGetDeviceCount(&nDev);
  if (ce != cudaSuccess || nDev == 0)
    return fail(VSLAMB200_ECUDA, std::string("no CUDA device available (") + cudaGetErrorString(ce) +
                                   "); vslam_b200 has no CPU fallback");
  if (device < 0 || device >= nDev) return fail(VSLAMB200_EINVAL, "device index out of range");
  CK(cudaSetDevice(device));
  auto c = new vslamb200_ctx();
  c->device = device;
  c->cfg = *cfg;
  std::memset(&c->stats, 0, sizeof(c->stats));
  cudaError_t e1 = cudaStreamCreateWithFlags(&c->own, cudaStreamNonBlocking);
  cudaError_t e2 = cudaSuccess;
  cudaError_t e3 = cudaEventCreate(&c->ev[0]);
  cudaError_t e4 = cudaEventCreate(&c->ev[1]);
  if (e4 == cudaSuccess) e4 = cudaEventCreate(&c->ev[2]);
  if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess || e4 != cudaSuccess) {
    delete c;
    return fail(VSLAMB200_ECUDA, "context resources could not be created");
  }
  c->stream = c->own;
  bool ok = cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&c->copyDone, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&c->computeMark, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&c->prepStream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&c->u16Done, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; i < vslamb200_ctx::kSlotRing; i++)
    ok = ok && cudaEventCreateWithFlags(&c->slotsEv[i], cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&c->forkEv, cudaEventDisableTiming) == cudaSuccess;
  for (int q = 0; q < vslamb200_ctx::kMaxParts; q++) {
    ok = ok && cudaStreamCreateWithFlags(&c->part[q], cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&c->joinEv[q], cudaEventDisableTiming) == cudaSuccess;
  }
  if (!ok) {
    vslamb200_destroy(c);
    return fail(VSLAMB200_ECUDA, "context resources could not be created");
  }
  *out = c;
  return 0;
}

void vslamb200_destroy(vslamb200_ctx * c)
{
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->copyStream) cudaStreamSynchronize(c->copyStream);
  if (c->prepStream) cudaStreamSynchronize(c->prepStream);
  for (auto & a : c->arenas) {
    if (a.ready) cudaEventDestroy(a.ready);
    if (a.ptr) cudaFree(a.ptr);
  }
  DevBuf * bufs[] = {&c->dFrames, &c->stageGray, &c->stageDepth, &c->stageU16, &c->slotsBuf, &c->jobs, &c->views,
                     &c->chunkCounts, &c->partials, &c->lists, &c->trace, &c->plane, &c->ext};
  for (auto b : bufs) b->release();
  if (c->copyStream) {
    cudaStreamSynchronize(c->copyStream);
    cudaStreamDestroy(c->copyStream);
  }
  if (c->prepStream) cudaStreamDestroy(c->prepStream);
  if (c->u16Done) cudaEventDestroy(c->u16Done);
  if (c->copyDone) cudaEventDestroy(c->copyDone);
  if (c->forkEv) cudaEventDestroy(c->forkEv);
  for (int q = 0; q < vslamb200_ctx::kMaxParts; q++) {
    if (c->part[q]) cudaStreamDestroy(c->part[q]);
    if (c->joinEv[q]) cudaEventDestroy(c->joinEv[q]);
  }
  if (c->computeMark) cudaEventDestroy(c->computeMark);
  for (int i = 0; i < vslamb200_ctx::kSlotRing; i++) {
    if (c->slotsEv[i]) cudaEventDestroy(c->slotsEv[i]);
    if (c->hSlots[i]) cudaFreeHost(c->hSlots[i]);
    if (i == 0 && c->hStage) cudaFreeHost(c->hStage);
    c->dSlots[i].release();
  }
  if (c->ev[0]) cudaEventDestroy(c->ev[0]);
  if (c->ev[1]) cudaEventDestroy(c->ev[1]);
  if (c->ev[2]) cudaEventDestroy(c->ev[2]);
  if (c->own) cudaStreamDestroy(c->own);
  delete c;
}

int vslamb200_set_config(vslamb200_ctx * c, const vslamb200_config * cfg)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  if (int e = check_cfg(cfg)) return e;
  c->cfg = *cfg;
  return 0;
}

int vslamb200_set_stream(vslamb200_ctx * c, void * s)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  CK(cudaStreamSynchronize(c->stream));
  c->stream = s ? (cudaStream_t)s : c->own;
  return 0;
}

int vslamb200_synchronize(vslamb200_ctx * c)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->copyStream));
  CK(cudaStreamSynchronize(c->prepStream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int vslamb200_frames_upload(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * gray, size_t grayStride,
  const void * depth, int depthType, size_t depthStride, int layout, const double * cams, int nLevels)
{
  if (!c || !gray || !depth) return fail(VSLAMB200_EINVAL, "frames_upload: null argument");
  if (depthType != VSLAMB200_DEPTH_F32 && depthType != VSLAMB200_DEPTH_F64)
    return fail(VSLAMB200_EINVAL, "frames_upload: bad depth_type");
  if (layout != VSLAMB200_ROW_MAJOR && layout != VSLAMB200_COL_MAJOR)
    return fail(VSLAMB200_EINVAL, "frames_upload: bad layout");
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots)) return e;
  const size_t npx = (size_t)W * H;
  const size_t dsz = depthType == VSLAMB200_DEPTH_F64 ? 8 : 4;
  if (!grayStride) grayStride = npx;
  if (!depthStride) depthStride = npx * dsz;
  // frames of one call live in one arena with a constant stride
  const vb::FrameDev & f0 = c->frames[slots[0]].dev;
  const size_t frameStride = n > 1 ? (size_t)((uint8_t *)c->frames[slots[1]].dev.I[0] - (uint8_t *)f0.I[0]) : npx * 4;  // one frame: any pitch >= the widest row copied
  const bool direct = layout == VSLAMB200_ROW_MAJOR && depthType == VSLAMB200_DEPTH_F32;
  if (direct) {
    CK(cudaMemcpy2DAsync(f0.I[0], frameStride, gray, grayStride, npx, n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpy2DAsync(f0.Z[0], frameStride, depth, depthStride, npx * 4, n, cudaMemcpyHostToDevice, c->stream));
  } else {
    const int sub = 32;  // frames per staging round
    CK(c->stageGray.ensure(npx * sub));
    CK(c->stageDepth.ensure(npx * dsz * sub));
    CK(c->slotsBuf.ensure((size_t)n * sizeof(int)));
    for (int b = 0; b < n; b += sub) {
      const int m = std::min(sub, n - b);
      CK(cudaMemcpy2DAsync(c->stageGray.p, npx, gray + grayStride * b, grayStride, npx, m, cudaMemcpyHostToDevice, c->stream));
      CK(cudaMemcpy2DAsync(c->stageDepth.p, npx * dsz, (const uint8_t *)depth + depthStride * b, depthStride, npx * dsz,
                           m, cudaMemcpyHostToDevice, c->stream));
      CK(cudaMemcpyAsync(c->slotsBuf.p, slots.data() + b, m * sizeof(int), cudaMemcpyHostToDevice, c->stream));
      vb::launch_ingest(c->stageGray.as<uint8_t>(), c->stageDepth.p, depthType, layout, W, H,
                        c->dFrames.as<vb::FrameDev>(), c->slotsBuf.as<int>(), m, c->stream);
      c->stats.kernel_launches++;
      CK(cudaStreamSynchronize(c->stream));
    }
  }
  c->stats.h2d_bytes += (uint64_t)n * npx * (1 + dsz);
  return preprocess(c, slots, nLevels, true, c->stream);
}

int vslamb200_frames_upload_async(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * gray, size_t grayStride,
  const void * depth, int depthType, size_t depthStride, int layout, const double * cams, int nLevels)
{
  if (!c || !gray || !depth) return fail(VSLAMB200_EINVAL, "frames_upload_async: null argument");
  if (layout != VSLAMB200_ROW_MAJOR || depthType != VSLAMB200_DEPTH_F32)  // the transposing ingest path stages: synchronous
    return vslamb200_frames_upload(c, n, ids, W, H, gray, grayStride, depth, depthType, depthStride, layout, cams, nLevels);
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots)) return e;
  const size_t npx = (size_t)W * H;
  if (!grayStride) grayStride = npx;
  if (!depthStride) depthStride = npx * 4;
  const vb::FrameDev & f0 = c->frames[slots[0]].dev;
  const size_t frameStride = n > 1 ? (size_t)((uint8_t *)c->frames[slots[1]].dev.I[0] - (uint8_t *)f0.I[0]) : npx * 4;  // one frame: any pitch >= the widest row copied
  // the copy may land in a recycled arena: order it after the compute work issued so far, not after what follows
  CK(cudaEventRecord(c->computeMark, c->stream));
  CK(cudaStreamWaitEvent(c->copyStream, c->computeMark, 0));
  CK(cudaMemcpy2DAsync(f0.I[0], frameStride, gray, grayStride, npx, n, cudaMemcpyHostToDevice, c->copyStream));
  CK(cudaMemcpy2DAsync(f0.Z[0], frameStride, depth, depthStride, npx * 4, n, cudaMemcpyHostToDevice, c->copyStream));
  c->stats.h2d_bytes += (uint64_t)n * npx * 5;
  // pyramids and gradients run on the preparation stream behind the copies: neither the compute stream (which does not
  // depend on this upload) nor the next transfer is held up by them; whoever reads these frames first waits for the
  // arena's event
  CK(cudaEventRecord(c->copyDone, c->copyStream));
  CK(cudaStreamWaitEvent(c->prepStream, c->copyDone, 0));
  if (int e = preprocess(c, slots, nLevels, false, c->prepStream)) return e;
  Arena & a = c->arenas[c->frames[slots[0]].arena];
  if (!a.ready) CK(cudaEventCreateWithFlags(&a.ready, cudaEventDisableTiming));
  CK(cudaEventRecord(a.ready, c->prepStream));
  a.pending = true;
  return 0;
}

int vslamb200_frames_upload_u16_async(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * gray, const uint16_t * depth, double depthDivisor,
  const double * cams, int nLevels)
{
  if (!c || !gray || !depth) return fail(VSLAMB200_EINVAL, "frames_upload_u16_async: null argument");
  if (!(depthDivisor > 0.0) || !std::isfinite(depthDivisor)) return fail(VSLAMB200_EINVAL, "frames_upload_u16_async: bad depth_divisor");
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots)) return e;
  const size_t npx = (size_t)W * H;
  const vb::FrameDev & f0 = c->frames[slots[0]].dev;
  const size_t frameStride = n > 1 ? (size_t)((uint8_t *)c->frames[slots[1]].dev.I[0] - (uint8_t *)f0.I[0]) : npx * 4;
  // the 16-bit planes land in a staging buffer of the context (grown once), the grey planes go straight into the frames;
  // conversion, pyramids and gradients run on the preparation stream behind the copies, so the next upload's transfer
  // starts as soon as this one's ends (the staging buffer is free again once the conversion kernel has run: u16Done)
  if (npx * 2 * (size_t)n > c->stageU16.cap) {
    CK(cudaStreamSynchronize(c->copyStream));
    CK(cudaStreamSynchronize(c->prepStream));
    CK(c->stageU16.ensure(npx * 2 * (size_t)n));
  }
  CK(cudaEventRecord(c->computeMark, c->stream));
  CK(cudaStreamWaitEvent(c->copyStream, c->computeMark, 0));
  CK(cudaMemcpy2DAsync(f0.I[0], frameStride, gray, npx, npx, n, cudaMemcpyHostToDevice, c->copyStream));
  if (c->u16Pending) CK(cudaStreamWaitEvent(c->copyStream, c->u16Done, 0));
  CK(cudaMemcpyAsync(c->stageU16.p, depth, npx * 2 * (size_t)n, cudaMemcpyHostToDevice, c->copyStream));
  c->stats.h2d_bytes += (uint64_t)n * npx * 3;
  CK(cudaEventRecord(c->copyDone, c->copyStream));
  CK(cudaStreamWaitEvent(c->prepStream, c->copyDone, 0));
  if (int e = preprocess(c, slots, nLevels, false, c->prepStream, c->stageU16.as<uint16_t>(), depthDivisor, c->u16Done)) return e;
  c->u16Pending = true;
  Arena & a = c->arenas[c->frames[slots[0]].arena];
  if (!a.ready) CK(cudaEventCreateWithFlags(&a.ready, cudaEventDisableTiming));
  CK(cudaEventRecord(a.ready, c->prepStream));
  a.pending = true;
  return 0;
}

int vslamb200_frames_upload_device(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * dGray, const float * dDepth,
  const double * cams, int nLevels)
{
  if (!c || !dGray || !dDepth) return fail(VSLAMB200_EINVAL, "frames_upload_device: null argument");
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots)) return e;
  const size_t npx = (size_t)W * H;
  const vb::FrameDev & f0 = c->frames[slots[0]].dev;
  const size_t frameStride = n > 1 ? (size_t)((uint8_t *)c->frames[slots[1]].dev.I[0] - (uint8_t *)f0.I[0]) : npx * 4;  // one frame: any pitch >= the widest row copied
  CK(cudaMemcpy2DAsync(f0.I[0], frameStride, dGray, npx, npx, n, cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaMemcpy2DAsync(f0.Z[0], frameStride, dDepth, npx * 4, npx * 4, n, cudaMemcpyDeviceToDevice, c->stream));
  return preprocess(c, slots, nLevels, true, c->stream);
}

int vslamb200_frames_attach_device(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * dGray, const float * dDepth,
  const double * cams, int nLevels)
{
  if (!c || !dGray || !dDepth) return fail(VSLAMB200_EINVAL, "frames_attach_device: null argument");
  if ((((uintptr_t)dGray) | ((uintptr_t)dDepth)) & 15)
    return fail(VSLAMB200_EINVAL, "frames_attach_device: the device buffers must be 16-byte aligned");
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots, dGray, dDepth)) return e;
  return preprocess(c, slots, nLevels, true, c->stream);
}

int vslamb200_frames_ingest_rgbd(
  vslamb200_ctx * c, int n, const uint64_t * ids, int W, int H, const uint8_t * rgb, int channelOrder, const void * depth,
  int depthType, double depthDivisor, const double * cams, int nLevels)
{
  if (!c || !rgb || !depth) return fail(VSLAMB200_EINVAL, "frames_ingest_rgbd: null argument");
  if (depthType != VSLAMB200_DEPTH_F32 && depthType != VSLAMB200_DEPTH_F64 && depthType != VSLAMB200_DEPTH_U16)
    return fail(VSLAMB200_EINVAL, "frames_ingest_rgbd: bad depth_type");
  if (channelOrder != VSLAMB200_RGB && channelOrder != VSLAMB200_BGR)
    return fail(VSLAMB200_EINVAL, "frames_ingest_rgbd: bad channel_order");
  if (!(depthDivisor > 0.0) || !std::isfinite(depthDivisor)) return fail(VSLAMB200_EINVAL, "frames_ingest_rgbd: bad depth_divisor");
  CK(cudaSetDevice(c->device));
  std::vector<int> slots;
  if (int e = make_frames(c, n, ids, W, H, cams, nLevels, slots)) return e;
  const size_t npx = (size_t)W * H;
  const size_t dsz = depthType == VSLAMB200_DEPTH_F64 ? 8 : depthType == VSLAMB200_DEPTH_U16 ? 2 : 4;
  const int sub = 32;  // frames per staging round
  CK(c->stageGray.ensure(npx * 3 * sub));
  CK(c->stageDepth.ensure(npx * dsz * sub));
  CK(c->slotsBuf.ensure((size_t)n * sizeof(int)));
  for (int b = 0; b < n; b += sub) {
    const int m = std::min(sub, n - b);
    CK(cudaMemcpyAsync(c->stageGray.p, rgb + npx * 3 * b, npx * 3 * m, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->stageDepth.p, (const uint8_t *)depth + npx * dsz * b, npx * dsz * m, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->slotsBuf.p, slots.data() + b, m * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    vb::launch_ingest_rgbd(c->stageGray.as<uint8_t>(), c->stageDepth.p, depthType, channelOrder == VSLAMB200_BGR, depthDivisor, W, H,
                           c->dFrames.as<vb::FrameDev>(), c->slotsBuf.as<int>(), m, c->stream);
    c->stats.kernel_launches++;
    CK(cudaStreamSynchronize(c->stream));
  }
  c->stats.h2d_bytes += (uint64_t)n * npx * (3 + dsz);
  return preprocess(c, slots, nLevels, true, c->stream);
}

int vslamb200_frame_release(vslamb200_ctx * c, uint64_t id)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  auto it = c->idToSlot.find(id);
  if (it == c->idToSlot.end()) return fail(VSLAMB200_ENOFRAME, "frame_release: unknown frame id");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->copyStream));
  CK(cudaStreamSynchronize(c->prepStream));
  CK(cudaStreamSynchronize(c->stream));
  return release_slot(c, it->second);
}

int vslamb200_frame_release_all(vslamb200_ctx * c)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->copyStream));
  CK(cudaStreamSynchronize(c->prepStream));
  CK(cudaStreamSynchronize(c->stream));
  for (size_t i = 0; i < c->frames.size(); i++) release_slot(c, (int)i);
  trim_spares(c, 0);
  return 0;
}

int vslamb200_frame_dims(vslamb200_ctx * c, uint64_t id, int level, int * d)
{
  if (!c || !d) return fail(VSLAMB200_EINVAL, "frame_dims: null argument");
  auto it = c->idToSlot.find(id);
  if (it == c->idToSlot.end()) return fail(VSLAMB200_ENOFRAME, "frame_dims: unknown frame id");
  const vb::FrameDev & f = c->frames[it->second].dev;
  if (level < 0 || level >= f.nLevels) return fail(VSLAMB200_EINVAL, "frame_dims: no such level");
  d[0] = f.W[level];
  d[1] = f.H[level];
  d[2] = f.Wd[level];
  d[3] = f.Hd[level];
  return 0;
}

int vslamb200_frame_download(
  vslamb200_ctx * c, uint64_t id, int level, uint8_t * I, float * Z, float * Zsf, int16_t * gx, int16_t * gy)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  auto it = c->idToSlot.find(id);
  if (it == c->idToSlot.end()) return fail(VSLAMB200_ENOFRAME, "frame_download: unknown frame id");
  const vb::FrameDev & f = c->frames[it->second].dev;
  if (level < 0 || level >= f.nLevels) return fail(VSLAMB200_EINVAL, "frame_download: no such level");
  CK(cudaSetDevice(c->device));
  if (int e = wait_frame_ready(c, it->second)) return e;
  const size_t n = (size_t)f.W[level] * f.H[level], nd = (size_t)f.Wd[level] * f.Hd[level];
  if (I) CK(cudaMemcpyAsync(I, f.I[level], n, cudaMemcpyDeviceToHost, c->stream));
  if (Z) CK(cudaMemcpyAsync(Z, f.Z[level], nd * 4, cudaMemcpyDeviceToHost, c->stream));
  if (Zsf) CK(cudaMemcpyAsync(Zsf, f.Zsf[level], nd * 4, cudaMemcpyDeviceToHost, c->stream));
  if (gx) CK(cudaMemcpyAsync(gx, f.gx[level], n * 2, cudaMemcpyDeviceToHost, c->stream));
  if (gy) CK(cudaMemcpyAsync(gy, f.gy[level], n * 2, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// batched alignment driver
// ------------------------------------------------------------------------------------------------
namespace
{
vb::Pose to_pose(const double * p)
{
  vb::Pose o;
  for (int i = 0; i < 4; i++) o.q[i] = p[i];
  for (int i = 0; i < 3; i++) o.t[i] = p[4 + i];
  return o;
}

vb::SolverParams solver_params(const vslamb200_config & cfg)
{
  vb::SolverParams p;
  p.minGradient2 = cfg.min_gradient * cfg.min_gradient;  // SE3Alignment.cpp:54
  p.lossKind = cfg.loss_kind;
  p.huberC = cfg.huber_c;
  p.lossV = cfg.loss_v;
  p.scalerKind = cfg.loss_kind == VSLAMB200_LOSS_NONE ? VSLAMB200_SCALER_IDENTITY : cfg.scaler_kind;
  p.scalerV = cfg.scaler_v;
  p.minStepSize = cfg.min_step_size;
  p.maxIterations = cfg.max_iterations;
  p.includePrior = cfg.include_prior;
  p.solverKind = cfg.solver_kind;
  p.lmLambda0 = cfg.lm_lambda0;
  p.lmUp = cfg.lm_up;
  p.lmDown = cfg.lm_down;
  p.sampling = cfg.sampling;
  p.depthWeight = cfg.depth_weight;
  p.chi2Mode = cfg.chi2_mode;
  return p;
}

struct Problem
{
  std::vector<vb::JobState> jobs;
  std::vector<vb::ViewState> views;
  int W = 0, H = 0, nLevels = 0, maxViews = 1;
  bool uniformViews = true;  // every job has the same number of keyframes
  vb::FrameDev geo;
};

// validate the jobs, build host-side job / view tables
int build_problem(vslamb200_ctx * c, const vslamb200_job * jobs, int nJobs, Problem & P)
{
  if (!jobs || nJobs <= 0) return fail(VSLAMB200_EINVAL, "align: no jobs");
  P.jobs.resize(nJobs);
  bool first = true;
  for (int j = 0; j < nJobs; j++) {
    const vslamb200_job & in = jobs[j];
    if (in.n_from < 1 || in.n_from > vb::MAX_VIEWS || !in.from || !in.from_poses)
      return fail(VSLAMB200_EINVAL, "align: n_from out of range or null keyframe list");
    if (!in.multi && in.n_from != 1) return fail(VSLAMB200_EINVAL, "align: single-frame overload takes one keyframe");
    auto itTo = c->idToSlot.find(in.to);
    if (itTo == c->idToSlot.end()) return fail(VSLAMB200_ENOFRAME, "align: unknown target frame id");
    vb::JobState J;
    std::memset(&J, 0, sizeof(J));
    J.nViews = in.n_from;
    J.view0 = (int)P.views.size();
    J.multi = in.multi ? 1 : 0;
    J.toPose = to_pose(in.to_pose);
    J.pose = J.toPose;
    for (int i = 0; i < 36; i++) J.cov[i] = (i % 7 == 0) ? 1.0 : 0.0;
    const vb::FrameDev & ft = c->frames[itTo->second].dev;
    if (int e = wait_frame_ready(c, itTo->second)) return e;
    for (int k = 0; k < in.n_from; k++) {
      auto itF = c->idToSlot.find(in.from[k]);
      if (itF == c->idToSlot.end()) return fail(VSLAMB200_ENOFRAME, "align: unknown keyframe id");
      if (int e = wait_frame_ready(c, itF->second)) return e;
      const vb::FrameDev & ff = c->frames[itF->second].dev;
      if (first) {
        P.geo = ff;
        P.W = ff.W[0];
        P.H = ff.H[0];
        P.nLevels = ff.nLevels;
        first = false;
      }
      for (const vb::FrameDev * f : {&ff, &ft}) {
        if (f->nLevels != P.nLevels || f->W[0] != P.W || f->H[0] != P.H)
          return fail(VSLAMB200_EINVAL, "align: all frames of one call must share size and pyramid depth");
      }
      vb::ViewState V;
      std::memset(&V, 0, sizeof(V));
      V.job = j;
      V.fromSlot = itF->second;
      V.toSlot = itTo->second;
      V.poseRef = to_pose(in.from_poses + 7 * k);
      P.views.push_back(V);
    }
    J.nLevels = P.nLevels;
    P.maxViews = std::max(P.maxViews, in.n_from);
    if (in.n_from != jobs[0].n_from) P.uniformViews = false;
    P.jobs[j] = J;
  }
  // the reference indexes the point cloud (depth dims) with the intensity width; at odd sizes the
  // two pyramids disagree and the reference is ill-defined
  for (int l = 0; l < P.nLevels; l++)
    if (P.geo.W[l] != P.geo.Wd[l] || P.geo.H[l] != P.geo.Hd[l])
      return fail(VSLAMB200_EDIMS, "align: intensity and depth pyramid sizes disagree (image size not divisible by 2^(levels-1))");
  return 0;
}

// cluster size of the dense level kernel: enough CTAs per job that a slice of the interest list is a few ten thousand
// points, never more than the portable maximum of 8
int pick_cluster(const vslamb200_config & cfg, int npx, int nJobs)
{
  const int o = cfg.cluster_size;  // tuning override
  if (o == 1 || o == 2 || o == 4 || o == 8 || o == 16) return o;
  int c = npx > 160000 ? 8 : npx > 40000 ? 4 : npx > 12000 ? 2 : 1;
  // Large launches (many waves of clusters): half the size on the two fine levels - the serial sections of an evaluation
  // cost CTA time in proportion to the cluster size.  Measured on 4096 pairs (two launches of 2048 jobs per level):
  // level 0 49.4 ms with 4 CTAs against 50.6 with 8 (53.2 with 2, 67.7 with 1), level 1 18.2 ms with 2 against 19.4-20.0
  // with 4 (21.0 with 1), level 2 4.9 ms with 2 against 5.3 with 1 and 6.3 with 4.
  if (nJobs >= 1024) c = npx > 160000 ? 4 : npx > 12000 ? 2 : 1;
  // Latency mode: a handful of jobs cannot fill the device, so each job spreads over as many SMs as a cluster can
  // span (16 CTAs is the non-portable cluster maximum of sm_100; the launcher opts in).  Measured on one 640x480
  // pair, 4 levels: 3.43 ms with the throughput sizes 8/4/2/1, 2.88 ms with 8 everywhere, 2.13 ms with 16 everywhere.
  // Throughput batches keep the smaller clusters: there the per-CTA work amortises the serial sections better.
  // (the same for any batch whose 16-CTA clusters all fit on the device at once: four 128-thread CTAs per SM)
  if (nJobs <= 8 || nJobs * 16 <= 4 * 148) c = npx > 12000 ? 16 : npx > 3000 ? 8 : 4;
  // up to two such waves still gain on the two fine levels (64 pairs: 1.81 + 1.93 ms instead of 2.14 + 2.44 ms; the coarse
  // levels and 128 pairs lose)
  else if (nJobs * 16 <= 8 * 148 && npx > 40000) c = 16;
  // Small batches (what one GPU of a strong-scaling run gets): a launch ends with its slowest job, whose Gauss-Newton steps
  // run one after the other, so the middle levels take clusters twice as large.  Measured on 512 pairs: level 1 4.7 ms
  // instead of 6.3 ms, level 2 1.9 instead of 2.2; on 2048 pairs and more the smaller clusters win (10.4 against 12.1 ms).
  else if (nJobs <= 640 && npx > 12000 && c < 8) c *= 2;
  return c;
}

// The moment-based level kernel (gn_mom.cu) keeps per-residual-magnitude moments of one keyframe per CTA: it needs a
// weight that is a function of |r| (scalers without offset), the same number K <= 8 of keyframes in every job of the
// launch (the cluster's CTAs are split among them) and list indices that fit 20 bits.
bool use_moments(const vslamb200_config & cfg, const Problem & P, const vb::SolverParams & sp, int npx)
{
  if (cfg.level_kernel == VSLAMB200_KERNEL_DENSE) return false;
  // measured: with one keyframe per job the dense kernel is faster on every level; with stacked
  // keyframes the moment kernel runs the views side by side and is 2.6 x faster (8 keyframes, 1280x720)
  if (cfg.level_kernel == VSLAMB200_KERNEL_AUTO && P.maxViews == 1) return false;
  if (!P.uniformViews || P.maxViews > 8 || npx > (1 << 20)) return false;
  return sp.scalerKind == VSLAMB200_SCALER_IDENTITY || sp.scalerKind == VSLAMB200_SCALER_TDIST;
}

// CTAs per keyframe of the moment kernel (256 threads each): a slice of a few ten thousand points per CTA; small
// batches spread every job over as many SMs as a cluster can span (latency mode, as above)
int pick_cpv(const vslamb200_config & cfg, int npx, int nJobs, int K)
{
  int c = npx > 160000 ? 4 : npx > 40000 ? 2 : 1;
  // few jobs: as long as 16-CTA clusters (the non-portable maximum) for all of them fit on the device at once (two
  // 256-thread CTAs per SM), a job spreads over 16 CTAs - 16 alignments against 8 key frames are 128 CTAs otherwise
  const bool few = nJobs <= 8 || nJobs * 16 <= 2 * 148;
  if (few) c = npx > 12000 ? 16 : npx > 3000 ? 8 : 4;
  const int o = cfg.cluster_size;
  if (o == 1 || o == 2 || o == 4 || o == 8 || o == 16) c = o;
  const int cap = few ? 16 : 8;
  while (c > 1 && c * K > cap) c >>= 1;
  return c;
}

// One pyramid level for the jobs [job0, job0 + nJobs) (their views are contiguous: [view0, view0 + nViews)), queued on
// `stream`.
int run_level(vslamb200_ctx * c, Problem & P, int level, const vb::SolverParams & sp, int maxEvals, int job0, int nJobs,
              cudaStream_t stream)
{
  const int view0 = P.jobs[job0].view0;
  const int nViews = (job0 + nJobs < (int)P.jobs.size() ? P.jobs[job0 + nJobs].view0 : (int)P.views.size()) - view0;
  const int npx = P.geo.W[level] * P.geo.H[level];
  const int listChunks = vb::list_chunks_for(npx);
  vb::JobState * dJobs = c->jobs.as<vb::JobState>();
  vb::ViewState * dViews = c->views.as<vb::ViewState>();
  const vb::FrameDev * table = c->dFrames.as<vb::FrameDev>();
  double * dTrace = c->traceEnabled ? c->trace.as<double>() : nullptr;
  const size_t traceJobStride = (size_t)P.nLevels * sp.maxIterations * vb::TRACE_STRIDE;

  if (c->timing) CK(cudaEventRecord(c->ev[0], stream));
  vb::launch_level_setup(table, dJobs, dViews, view0, nViews, level, stream);
  const bool moments = sp.sampling != VSLAMB200_SAMPLING_BILINEAR && use_moments(c->cfg, P, sp, npx);
  vb::launch_build_lists(table, dViews, dJobs, view0, nViews, level, npx, sp.minGradient2, c->chunkCounts.as<int>(),
                         listChunks, moments ? 1 : 0, stream);
  vb::launch_level_begin(dJobs, dViews, job0, nJobs, level, stream);
  if (c->timing) CK(cudaEventRecord(c->ev[1], stream));
  if (sp.sampling == VSLAMB200_SAMPLING_BILINEAR) {
    // extension (ext.cu): three launches per evaluation over all jobs; every fourth evaluation the host reads how many
    // jobs are still running and stops queueing when none is
    int * dActive = c->ext.as<int>();
    for (int it = 0; it < maxEvals; it++) {
      const bool check = (it & 3) == 3 || it + 1 == maxEvals;
      vb::launch_ext_iteration(dJobs, dViews, job0, nJobs, view0, nViews, npx, sp, level, it + 1 >= maxEvals, dTrace,
                               traceJobStride, check ? dActive : nullptr, stream);
      c->stats.kernel_launches += 3 + (check ? 1 : 0);
      if (check && it + 1 < maxEvals) {
        int active = 0;
        CK(cudaMemcpyAsync(&active, dActive, sizeof(int), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        if (active == 0) break;
      }
    }
  } else if (moments) {
    c->stats.moment_launches++;
    CK(vb::launch_gn_mom(dJobs, dViews, job0, nJobs, P.maxViews, pick_cpv(c->cfg, npx, nJobs, P.maxViews), sp, level, maxEvals,
                         dTrace, traceJobStride, c->partials.as<double>(), stream));
  } else {
    c->stats.dense_launches++;
    CK(vb::launch_gn_level(dJobs, dViews, job0, nJobs, P.maxViews, pick_cluster(c->cfg, npx, nJobs), sp, level, maxEvals, dTrace,
                           traceJobStride, c->partials.as<double>(), stream));
  }
  if (c->timing) CK(cudaEventRecord(c->ev[2], stream));
  // level setup, list build, level begin, and the level kernel (the bilinear extension counted its own above)
  c->stats.kernel_launches += sp.sampling == VSLAMB200_SAMPLING_BILINEAR ? 3 : 4;
  CK(cudaGetLastError());
  if (c->timing) {
    CK(cudaStreamSynchronize(stream));
    float a = 0, b = 0;
    CK(cudaEventElapsedTime(&a, c->ev[0], c->ev[1]));
    CK(cudaEventElapsedTime(&b, c->ev[1], c->ev[2]));
    c->stats.ms_lists += a;
    c->stats.ms_gn_level[level] += b;
  }
  return 0;
}

int prepare_workspace(vslamb200_ctx * c, Problem & P, const vb::SolverParams & sp)
{
  const int nJobs = (int)P.jobs.size(), nViews = (int)P.views.size();
  const size_t npx0 = (size_t)P.W * P.H;
  const size_t listStride = align_up(npx0, vb::LIST_TILE);
  // per view and candidate pixel: list tile 8 B | record 16 B | residual code 2 B | sort permutation 4 B = 30 B
  const size_t tileBytes = listStride / vb::LIST_TILE * vb::LIST_TILE_BYTES;
  const size_t recBytes = listStride * sizeof(vb::ListRec);
  const size_t viewBytes = tileBytes + recBytes + listStride * 2 + listStride * 4 + 256;
  CK(c->lists.ensure(viewBytes * nViews));
  CK(c->chunkCounts.ensure((size_t)nViews * vb::list_chunks_for((int)npx0) * sizeof(int)));
  CK(c->partials.ensure((size_t)nViews * vb::NE_TERMS * sizeof(double)));
  uint8_t * base = c->lists.as<uint8_t>();
  for (int v = 0; v < nViews; v++) {
    vb::ViewState & V = P.views[v];
    uint8_t * b = base + (size_t)v * viewBytes;
    V.tiles = b;
    V.rec = (vb::ListRec *)(b + tileBytes);
    V.r = (int16_t *)(b + tileBytes + recBytes);
    V.perm = (uint32_t *)(b + tileBytes + recBytes + listStride * 2);
  }
  if (sp.sampling == VSLAMB200_SAMPLING_BILINEAR) {
    // bilinear extension: per view 16 B per candidate pixel (r_I, r_Z in fp64) + 28 partial sums per 2048 list entries
    const size_t chunks = (npx0 + vb::EXT_CHUNK - 1) / vb::EXT_CHUNK;
    const size_t rBytes = listStride * sizeof(double2), partBytes = align_up(chunks * vb::NE_TERMS * sizeof(double), 256);
    CK(c->ext.ensure((rBytes + partBytes) * nViews + 256));
    uint8_t * eb = c->ext.as<uint8_t>() + 256;  // the first 256 bytes hold the active-job counter
    for (int v = 0; v < nViews; v++) {
      P.views[v].extR = (double2 *)(eb + (size_t)v * (rBytes + partBytes));
      P.views[v].extPart = (double *)(eb + (size_t)v * (rBytes + partBytes) + rBytes);
    }
  }
  CK(c->jobs.ensure(nJobs * sizeof(vb::JobState)));
  CK(c->views.ensure(nViews * sizeof(vb::ViewState)));
  {
    // every alignment call ends with a stream synchronisation, so the staging buffer is free here
    const size_t jb = align_up(nJobs * sizeof(vb::JobState), 256), vbytes = nViews * sizeof(vb::ViewState);
    if (jb + vbytes > c->hStageCap) {
      if (c->hStage) cudaFreeHost(c->hStage);
      c->hStage = nullptr;
      c->hStageCap = 0;
      const size_t want = (jb + vbytes) * 3 / 2 + 4096;
      CK(cudaHostAlloc(&c->hStage, want, cudaHostAllocMapped | cudaHostAllocPortable));
      c->hStageCap = want;
    }
    std::memcpy(c->hStage, P.jobs.data(), nJobs * sizeof(vb::JobState));
    std::memcpy((uint8_t *)c->hStage + jb, P.views.data(), vbytes);
    vb::launch_stage_copy(c->jobs.p, c->hStage, nJobs * sizeof(vb::JobState), c->stream);
    vb::launch_stage_copy(c->views.p, (uint8_t *)c->hStage + jb, vbytes, c->stream);
    c->stats.kernel_launches += 2;
  }
  if (c->traceEnabled) {
    const size_t bytes = (size_t)nJobs * P.nLevels * sp.maxIterations * vb::TRACE_STRIDE * sizeof(double);
    CK(c->trace.ensure(bytes));
    CK(cudaMemsetAsync(c->trace.p, 0, bytes, c->stream));
    c->traceJobs = nJobs;
    c->traceLevels = P.nLevels;
    c->traceMaxIter = sp.maxIterations;
  }
  return 0;
}
}  // namespace

extern "C" {

int vslamb200_align(vslamb200_ctx * c, const vslamb200_job * jobs, int nJobs, vslamb200_result * out)
{
  if (!c || !out) return fail(VSLAMB200_EINVAL, "align: null argument");
  CK(cudaSetDevice(c->device));
  Problem P;
  if (int e = build_problem(c, jobs, nJobs, P)) return e;
  const vb::SolverParams sp = solver_params(c->cfg);
  if (int e = prepare_workspace(c, P, sp)) return e;
  // Jobs stop after different numbers of iterations, so every level kernel ends with a tail of a few long-running
  // clusters.  A large batch is therefore cut into two halves that walk down the pyramid on their own streams: one
  // half's tail is filled by the other half's next kernel.  (Timing mode keeps one stream: it measures whole launches.)
  int nParts = (!c->timing && nJobs >= 2 * kMinJobsPerPart && sp.sampling == VSLAMB200_SAMPLING_NEAREST) ? 2 : 1;
  if (c->cfg.stream_parts > 0) nParts = std::min(vslamb200_ctx::kMaxParts, c->cfg.stream_parts);
  nParts = std::min(nParts, nJobs);
  if (sp.sampling != VSLAMB200_SAMPLING_NEAREST) nParts = 1;  // the extension's host loop reads one active-job counter
  if (nParts == 1) {
    for (int level = P.nLevels - 1; level >= 0; level--)
      if (int e = run_level(c, P, level, sp, sp.maxIterations + 1, 0, nJobs, c->stream)) return e;
  } else {
    CK(cudaEventRecord(c->forkEv, c->stream));
    int cut[vslamb200_ctx::kMaxParts + 1];
    for (int q = 0; q <= nParts; q++) cut[q] = (int)((long long)nJobs * q / nParts);
    for (int q = 0; q < nParts; q++) CK(cudaStreamWaitEvent(c->part[q], c->forkEv, 0));
    for (int level = P.nLevels - 1; level >= 0; level--)
      for (int q = 0; q < nParts; q++)
        if (int e = run_level(c, P, level, sp, sp.maxIterations + 1, cut[q], cut[q + 1] - cut[q], c->part[q])) return e;
    for (int q = 0; q < nParts; q++) {
      CK(cudaEventRecord(c->joinEv[q], c->part[q]));
      CK(cudaStreamWaitEvent(c->stream, c->joinEv[q], 0));
    }
  }
  CK(cudaMemcpyAsync(P.jobs.data(), c->jobs.p, nJobs * sizeof(vb::JobState), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->stats.d2h_bytes += (uint64_t)nJobs * sizeof(vb::JobState);
  for (int j = 0; j < nJobs; j++) {
    const vb::JobState & J = P.jobs[j];
    vslamb200_result & r = out[j];
    for (int i = 0; i < 4; i++) r.pose[i] = J.pose.q[i];
    for (int i = 0; i < 3; i++) r.pose[4 + i] = J.pose.t[i];
    std::memcpy(r.cov, J.cov, sizeof(r.cov));
    r.status = J.status;
    for (int l = 0; l < VSLAMB200_MAX_LEVELS; l++) {
      r.iters[l] = l < P.nLevels ? J.iters[l] : 0;
      r.stop[l] = l < P.nLevels ? J.stop[l] : 0;
      if (l < P.nLevels) {
        const uint64_t px = (uint64_t)J.evals[l] * J.nViews * P.geo.W[l] * P.geo.H[l];
        c->stats.pixel_iterations += px;
        c->stats.pixel_iterations_level[l] += px;
        for (int k = 0; k < 10; k++) c->stats.phase_clocks[l][k] += (uint64_t)J.clk[l][k];
        c->stats.gn_iterations += J.evals[l];
      }
    }
    c->stats.point_iterations += J.pointIters;
  }
  return 0;
}

int vslamb200_set_trace(vslamb200_ctx * c, int enable)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  c->traceEnabled = enable ? 1 : 0;
  return 0;
}

int vslamb200_get_trace(vslamb200_ctx * c, int job, double * out)
{
  if (!c || !out) return fail(VSLAMB200_EINVAL, "get_trace: null argument");
  if (!c->trace.p || job < 0 || job >= c->traceJobs) return fail(VSLAMB200_EINVAL, "get_trace: no trace for that job");
  CK(cudaSetDevice(c->device));
  const size_t n = (size_t)c->traceLevels * c->traceMaxIter * vb::TRACE_STRIDE;
  CK(cudaMemcpy(out, c->trace.as<double>() + n * job, n * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

int vslamb200_evaluate(
  vslamb200_ctx * c, uint64_t from, uint64_t to, int level, const double * se3, int multiMask, int16_t * rcode,
  double * ne64)
{
  if (!c || !se3) return fail(VSLAMB200_EINVAL, "evaluate: null argument");
  CK(cudaSetDevice(c->device));
  const double ident[7] = {0, 0, 0, 1, 0, 0, 0};
  vslamb200_job job;
  job.to = to;
  std::memcpy(job.to_pose, se3, sizeof(job.to_pose));
  job.n_from = 1;
  job.multi = multiMask ? 1 : 0;
  job.from = &from;
  job.from_poses = ident;
  Problem P;
  if (int e = build_problem(c, &job, 1, P)) return e;
  if (level < 0 || level >= P.nLevels) return fail(VSLAMB200_EINVAL, "evaluate: no such level");
  if (c->cfg.sampling != VSLAMB200_SAMPLING_NEAREST)
    return fail(VSLAMB200_EINVAL, "evaluate: single evaluations are defined for nearest-pixel sampling only");
  vb::SolverParams sp = solver_params(c->cfg);
  sp.includePrior = 0;
  const int savedTrace = c->traceEnabled;
  c->traceEnabled = 0;
  int e = prepare_workspace(c, P, sp);
  c->traceEnabled = savedTrace;
  if (e) return e;
  // one computeNormalEquations with the warp forced to exactly the caller's transform
  const int nViews = 1, nJobs = 1;
  const int npx = P.geo.W[level] * P.geo.H[level];
  vb::JobState * dJobs = c->jobs.as<vb::JobState>();
  vb::ViewState * dViews = c->views.as<vb::ViewState>();
  const vb::FrameDev * table = c->dFrames.as<vb::FrameDev>();
  vb::launch_level_setup(table, dJobs, dViews, 0, nViews, level, c->stream);
  vb::launch_build_lists(table, dViews, dJobs, 0, nViews, level, npx, sp.minGradient2, c->chunkCounts.as<int>(),
                         vb::list_chunks_for(npx), use_moments(c->cfg, P, sp, npx) ? 1 : 0, c->stream);
  vb::launch_level_begin(dJobs, dViews, 0, nJobs, level, c->stream);
  vb::ViewState V;
  CK(cudaMemcpyAsync(&V, dViews, sizeof(V), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  V.se3 = to_pose(se3);
  vb::view_set_warp(V);
  vb::view_set_x(V);
  CK(cudaMemcpyAsync(dViews, &V, sizeof(V), cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->partials.p, 0, vb::NE_TERMS * sizeof(double), c->stream));
  if (use_moments(c->cfg, P, sp, npx)) {
    c->stats.moment_launches++;
    CK(vb::launch_gn_mom(dJobs, dViews, 0, nJobs, 1, pick_cpv(c->cfg, npx, 1, 1), sp, level, 1, nullptr, 0, c->partials.as<double>(),
                         c->stream));
  } else {
    c->stats.dense_launches++;
    CK(vb::launch_gn_level(dJobs, dViews, 0, nJobs, 1, pick_cluster(c->cfg, npx, 1), sp, level, 1, nullptr, 0,
                           c->partials.as<double>(), c->stream));
  }
  c->stats.kernel_launches += 4;
  if (rcode) {
    CK(c->plane.ensure((size_t)npx * 2));
    vb::launch_scatter_rcode(dViews, 0, c->plane.as<int16_t>(), c->stream);
    CK(cudaMemcpyAsync(rcode, c->plane.p, (size_t)npx * 2, cudaMemcpyDeviceToHost, c->stream));
    c->stats.kernel_launches += 2;
  }
  double sums[vb::NE_TERMS];
  CK(cudaMemcpyAsync(sums, c->partials.p, sizeof(sums), cudaMemcpyDeviceToHost, c->stream));
  vb::ViewState V2;
  CK(cudaMemcpyAsync(&V2, dViews, sizeof(V2), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  if (ne64) {
    double s[vb::NE_TERMS];
    std::memcpy(s, sums, sizeof(s));
    const double n = (double)V.nIp;
    if (V.nIp > 1)
      for (double & x : s) x /= n;
    std::memset(ne64, 0, 64 * sizeof(double));
    ne64[0] = n;
    ne64[1] = s[27];
    ne64[2] = V2.scale;
    ne64[3] = V2.offset;
    for (int i = 0; i < 6; i++) ne64[4 + i] = s[21 + i];
    int idx = 0;
    for (int i = 0; i < 6; i++)
      for (int j = i; j < 6; j++) {
        ne64[10 + i * 6 + j] = s[idx];
        ne64[10 + j * 6 + i] = s[idx++];
      }
  }
  return 0;
}

int vslamb200_get_stats(vslamb200_ctx * c, vslamb200_stats * out)
{
  if (!c || !out) return fail(VSLAMB200_EINVAL, "get_stats: null argument");
  *out = c->stats;
  return 0;
}
int vslamb200_reset_stats(vslamb200_ctx * c)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  std::memset(&c->stats, 0, sizeof(c->stats));
  return 0;
}
int vslamb200_set_timing(vslamb200_ctx * c, int enable)
{
  if (!c) return fail(VSLAMB200_EINVAL, "ctx is null");
  c->timing = enable ? 1 : 0;
  return 0;
}

}  // extern "C"
