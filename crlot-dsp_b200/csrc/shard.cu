// shard.cu -- the multi-GPU partitioning of the hot path behind the C-ABI (SURVEY 8e, 8b last row).
//
//   crlot_channel_shard / crlot_shard_plan    pure arithmetic (no device needed)
//   crlot_shard_group_*                       one NCCL communicator per process/device
//   crlot_halo_exchange                       the (nfft - hop)-sample neighbour exchange of a time shard
//   crlot_sharded_roundtrip_execute           halo exchange + K1 + (K2, K3 | fused K2+K3) at the shard's offsets
//
// Why this split is exact: frame k reads only x[k*hop, k*hop + nfft) (ref: dsp/frame/framer.cc:128-182) and
// output sample t sums only the frames with k*hop in (t - nfft, t], per channel, in ascending k
// (ref: dsp/ola/OLAAccumulator.cc:75-111,194-209).  Rank r owns the OUTPUT range [r*S, (r+1)*S); it receives
// nfft - hop input samples from BOTH neighbours and recomputes the ceil(nfft/hop) - 1 frames that start in its
// left halo, so every output sample sums the same terms in the same order as the one-GPU run: bit-identical.
//
// NCCL is bound at run time (dlopen of libnccl.so.2): the library itself has no link-time NCCL dependency, a
// process that already carries an NCCL (torch's) shares it, and single-GPU users never load it.  Only the
// stable C entry points are used; their prototypes are restated here (nccl.h: ncclSend/ncclRecv/...).
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include "../../include/crlot_cuda.h"

extern "C" void crlot_cuda_set_error_(const char* msg);

namespace {

int sfail(int code, const std::string& msg) {
  crlot_cuda_set_error_(msg.c_str());
  return code;
}

// ---- NCCL, bound lazily ---------------------------------------------------------------------------
typedef struct ncclComm* nccl_comm_t;
struct nccl_unique_id { char internal[CRLOT_NCCL_UNIQUE_ID_BYTES]; };
constexpr int kNcclFloat = 7;          // ncclFloat32 (nccl.h ncclDataType_t)

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(nccl_unique_id*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
  int (*CommInitAll)(nccl_comm_t*, int, const int*) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  std::string why;
};

NcclApi* nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {getenv("CRLOT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      if (!n || !*n) continue;
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
      api.why = dlerror();
    }
    if (!api.handle) return;
    auto sym = [&](const char* s) { return dlsym(api.handle, s); };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommInitAll || !api.CommDestroy || !api.GroupStart ||
        !api.GroupEnd || !api.Send || !api.Recv) {
      api.why = "libnccl lacks a required symbol";
      api.handle = nullptr;
    }
  });
  return &api;
}

int nccl_fail(int rc, const char* where) {
  NcclApi* a = nccl();
  std::string m = std::string(where) + ": NCCL error " + std::to_string(rc);
  if (a->GetErrorString) m += std::string(" (") + a->GetErrorString(rc) + ")";
  return sfail(CRLOT_ERR_RUNTIME, m);
}
#define NC(call)                                   \
  do {                                             \
    int rc__ = (call);                             \
    if (rc__ != 0) return nccl_fail(rc__, #call);  \
  } while (0)

int need_nccl() {
  NcclApi* a = nccl();
  if (!a->handle) return sfail(CRLOT_ERR_RUNTIME, "NCCL is not available: " + a->why);
  return CRLOT_OK;
}

struct Dev {
  int prev = -1;
  explicit Dev(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); }
  ~Dev() { if (prev >= 0) cudaSetDevice(prev); }
};

long long ceil_div(long long a, long long b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); }

// the arithmetic of one shard; *sh is fully written
void plan_shard(long long total_len, int nfft, int hop, int rank, int world, int mode, crlot_time_shard* sh) {
  std::memset(sh, 0, sizeof(*sh));
  sh->rank = rank; sh->world = world; sh->nfft = nfft; sh->hop = hop; sh->boundary_mode = mode; sh->total_len = total_len;
  const long long S = total_len / world;
  const long long F = crlot_frame_count(total_len, nfft, hop, mode);
  const long long n_out_total = F > 0 ? (F - 1) * hop + nfft : 0;
  sh->own_lo = rank * S; sh->own_hi = (rank + 1) * S;
  sh->t0 = std::min<long long>(rank * S, n_out_total);
  sh->t1 = rank == world - 1 ? n_out_total : std::min<long long>((rank + 1) * S, n_out_total);
  if (sh->t1 <= sh->t0 || F == 0) {          // nothing to produce here: an empty shard with empty halos
    sh->t1 = sh->t0;
    sh->in_lo = sh->in_hi = sh->own_lo;
    sh->own_hi = sh->own_lo;
    return;
  }
  sh->k_lo = std::max<long long>(0, ceil_div(sh->t0 - nfft + 1, hop));
  const long long k_hi = std::min<long long>(F - 1, (sh->t1 - 1) / hop);
  sh->n_frames = k_hi - sh->k_lo + 1;
  sh->in_lo = sh->k_lo * hop;
  sh->in_hi = std::min<long long>(total_len, k_hi * hop + nfft);
  const long long first_owned = std::max<long long>(sh->k_lo, ceil_div(sh->own_lo, hop));
  const long long last_owned = std::min<long long>(k_hi, (sh->own_hi - 1) / hop);
  sh->n_owned_frames = std::max<long long>(0, last_owned - first_owned + 1);
  sh->left_halo = sh->own_lo - sh->in_lo;
  sh->right_halo = std::max<long long>(0, sh->in_hi - sh->own_hi);
  sh->n_out = sh->t1 - sh->t0;
  sh->frame_t0 = sh->k_lo * hop - sh->t0;
}

}  // namespace

struct crlot_shard_group_s {
  nccl_comm_t comm = nullptr;
  int world = 1, rank = 0, device = 0;
  bool owned = false;
};

extern "C" {

int crlot_channel_shard(long long channels, int rank, int world, long long* first, long long* count) {
  if (!first || !count) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels < 0 || world <= 0 || rank < 0 || rank >= world) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad rank/world");
  const long long base = channels / world, rem = channels % world;
  *first = rank * base + std::min<long long>(rank, rem);
  *count = base + (rank < rem ? 1 : 0);
  return CRLOT_OK;
}

int crlot_shard_plan(long long total_len, int nfft, int hop, int rank, int world, int boundary_mode, crlot_time_shard* out) {
  if (!out) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (nfft <= 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "Frame size must be greater than 0");
  if (hop <= 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "Hop size must be greater than 0");
  if (world <= 0 || rank < 0 || rank >= world || total_len < 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad rank/world/length");
  if (boundary_mode != CRLOT_ZERO_PAD && boundary_mode != CRLOT_DROP) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad boundary mode");
  if (total_len % world != 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "signal length must split evenly over the ranks");
  const long long S = total_len / world;
  if (S % hop != 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "shard length must be a multiple of hop");
  plan_shard(total_len, nfft, hop, rank, world, boundary_mode, out);
  // what this rank sends = what the neighbour's own plan says it receives
  crlot_time_shard nb;
  if (rank > 0) { plan_shard(total_len, nfft, hop, rank - 1, world, boundary_mode, &nb); out->send_left = nb.right_halo; }
  if (rank < world - 1) { plan_shard(total_len, nfft, hop, rank + 1, world, boundary_mode, &nb); out->send_right = nb.left_halo; }
  // a halo must come from the adjacent rank only
  if (out->left_halo > S || out->right_halo > S || out->send_left > S || out->send_right > S)
    return sfail(CRLOT_ERR_INVALID_ARGUMENT, "shard length must be at least the halo (about nfft - hop samples)");
  return CRLOT_OK;
}

// ---- communicator ------------------------------------------------------------------------------------
int crlot_shard_group_unique_id(void* id_out) {
  if (!id_out) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  int rc = need_nccl();
  if (rc) return rc;
  nccl_unique_id id;
  NC(nccl()->GetUniqueId(&id));
  std::memcpy(id_out, &id, sizeof(id));
  return CRLOT_OK;
}

int crlot_shard_group_create(int world, int rank, const void* id_in, int device, crlot_shard_group_t* out) {
  if (!out) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (world <= 0 || rank < 0 || rank >= world) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad rank/world");
  auto* g = new (std::nothrow) crlot_shard_group_s;
  if (!g) return sfail(CRLOT_ERR_ALLOC, "host allocation failed");
  g->world = world; g->rank = rank; g->device = device;
  if (world > 1) {
    if (!id_in) { delete g; return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null unique id"); }
    int rc = need_nccl();
    if (rc) { delete g; return rc; }
    Dev d(device);
    nccl_unique_id id;
    std::memcpy(&id, id_in, sizeof(id));
    rc = nccl()->CommInitRank(&g->comm, world, id, rank);
    if (rc != 0) { delete g; return nccl_fail(rc, "ncclCommInitRank"); }
    g->owned = true;
  }
  *out = g;
  return CRLOT_OK;
}

int crlot_shard_group_create_all(int n_dev, const int* devices, crlot_shard_group_t* out) {
  if (!out || n_dev <= 0) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad arguments");
  nccl_comm_t comms[64];
  if (n_dev > 64) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "at most 64 devices");
  if (n_dev > 1) {
    int rc = need_nccl();
    if (rc) return rc;
    NC(nccl()->CommInitAll(comms, n_dev, devices));
  }
  for (int i = 0; i < n_dev; ++i) {
    auto* g = new (std::nothrow) crlot_shard_group_s;
    if (!g) return sfail(CRLOT_ERR_ALLOC, "host allocation failed");
    g->world = n_dev; g->rank = i; g->device = devices ? devices[i] : i;
    g->comm = n_dev > 1 ? comms[i] : nullptr; g->owned = n_dev > 1;
    out[i] = g;
  }
  return CRLOT_OK;
}

int crlot_shard_group_wrap(void* nccl_comm, int world, int rank, int device, crlot_shard_group_t* out) {
  if (!out || (world > 1 && !nccl_comm)) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (world <= 0 || rank < 0 || rank >= world) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "bad rank/world");
  if (world > 1) { int rc = need_nccl(); if (rc) return rc; }
  auto* g = new (std::nothrow) crlot_shard_group_s;
  if (!g) return sfail(CRLOT_ERR_ALLOC, "host allocation failed");
  g->comm = static_cast<nccl_comm_t>(nccl_comm); g->world = world; g->rank = rank; g->device = device; g->owned = false;
  *out = g;
  return CRLOT_OK;
}

int crlot_shard_group_destroy(crlot_shard_group_t g) {
  if (!g) return CRLOT_OK;
  if (g->owned && g->comm) { Dev d(g->device); nccl()->CommDestroy(g->comm); }
  delete g;
  return CRLOT_OK;
}

int crlot_shard_group_info(crlot_shard_group_t g, int* world, int* rank, int* device) {
  if (!g) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (world) *world = g->world;
  if (rank) *rank = g->rank;
  if (device) *device = g->device;
  return CRLOT_OK;
}

// ---- the neighbour exchange ---------------------------------------------------------------------------
// d_xbuf is laid out [left halo | owned | right halo]; four point-to-point operations at most, all inside ONE
// group so that the two directions of a boundary progress together; zero-length halves are skipped on both
// sides (sender and receiver derive the same counts from the same arithmetic).
int crlot_halo_exchange(crlot_shard_group_t g, const crlot_time_shard* sh, float* d_xbuf, void* stream) {
  if (!g || !sh) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (sh->world != g->world || sh->rank != g->rank) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "shard and group disagree on rank/world");
  if (g->world == 1) return CRLOT_OK;
  if (!d_xbuf) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  const long long own = sh->own_hi - sh->own_lo;
  if (sh->send_left > own || sh->send_right > own) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "halo exceeds the owned span");
  NcclApi* a = nccl();
  Dev d(g->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* owned = d_xbuf + sh->left_halo;
  NC(a->GroupStart());
  int rc = 0;
  if (g->rank > 0) {
    if (sh->send_left > 0 && !rc) rc = a->Send(owned, static_cast<size_t>(sh->send_left), kNcclFloat, g->rank - 1, g->comm, st);
    if (sh->left_halo > 0 && !rc) rc = a->Recv(d_xbuf, static_cast<size_t>(sh->left_halo), kNcclFloat, g->rank - 1, g->comm, st);
  }
  if (g->rank < g->world - 1) {
    if (sh->send_right > 0 && !rc) rc = a->Send(owned + own - sh->send_right, static_cast<size_t>(sh->send_right), kNcclFloat, g->rank + 1, g->comm, st);
    if (sh->right_halo > 0 && !rc) rc = a->Recv(owned + own, static_cast<size_t>(sh->right_halo), kNcclFloat, g->rank + 1, g->comm, st);
  }
  const int rc_end = a->GroupEnd();
  if (rc) return nccl_fail(rc, "ncclSend/ncclRecv");
  if (rc_end) return nccl_fail(rc_end, "ncclGroupEnd");
  return CRLOT_OK;
}

// ---- the sharded round trip -----------------------------------------------------------------------------
int crlot_sharded_roundtrip_execute(crlot_plan_t plan, crlot_shard_group_t g, const crlot_time_shard* sh, float* d_xbuf,
                                    void* d_spec, float* d_frames, float* d_out, void* stream) {
  if (!plan || !sh || !d_xbuf) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  crlot_plan_info info;
  int rc = crlot_stft_plan_info(plan, &info);
  if (rc) return rc;
  if (info.nfft != sh->nfft || info.hop != sh->hop) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "plan and shard disagree on frame size or hop");
  if (g) {
    rc = crlot_halo_exchange(g, sh, d_xbuf, stream);
    if (rc) return rc;
  } else if (sh->world != 1) {
    return sfail(CRLOT_ERR_INVALID_ARGUMENT, "a shard of a multi-rank split needs its group");
  }
  if (sh->n_frames == 0) return CRLOT_OK;
  if (!d_spec || !d_out) return sfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  rc = crlot_stft_execute(plan, d_xbuf, 0, sh->in_hi - sh->in_lo, 1, sh->n_frames, d_spec, stream);
  if (rc) return rc;
  if (!d_frames) return crlot_istft_ola_execute(plan, d_spec, sh->n_frames, sh->n_frames, 1, sh->t0, sh->frame_t0, d_out, 0, sh->n_out, stream);
  rc = crlot_istft_execute(plan, d_spec, sh->n_frames, d_frames, 1, stream);
  if (rc) return rc;
  return crlot_ola_accumulate(plan, d_frames, sh->n_frames, sh->n_frames, 1, 1, sh->t0, sh->frame_t0, d_out, 0, sh->n_out, stream);
}

}  // extern "C"
