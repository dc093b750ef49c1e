// K6: the one exchange step of the path -- the hand-over of per-rank sparse posterior
// fragments to the rank that runs the progressive alignment (SURVEY.md 8(e) "Collective").
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy a host process already
// loaded, e.g. torch's, else the system one), so libdafs_cuda has no link-time dependency
// on it and still loads where NCCL is absent; the comm entry points then fail with
// DAFS_CUDA_ENOTSUP.  One rank per process and device (the launch the contract uses).
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "pairhmm_plan.cuh"

namespace dafs {

struct NcclApi {
  void* so = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
};

static const NcclApi* nccl_api() {
  static const NcclApi* api = []() -> const NcclApi* {
    NcclApi* a = new NcclApi();
    a->so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);          // already in the process?
    if (!a->so) a->so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!a->so) a->so = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!a->so) return a;
    bool ok = true;
#define DAFS_SYM(field, name) do { a->field = reinterpret_cast<decltype(a->field)>(dlsym(a->so, name)); if (!a->field) ok = false; } while (0)
    DAFS_SYM(GetUniqueId, "ncclGetUniqueId"); DAFS_SYM(CommInitRank, "ncclCommInitRank");
    DAFS_SYM(CommDestroy, "ncclCommDestroy"); DAFS_SYM(AllGather, "ncclAllGather");
    DAFS_SYM(Send, "ncclSend"); DAFS_SYM(Recv, "ncclRecv");
    DAFS_SYM(GroupStart, "ncclGroupStart"); DAFS_SYM(GroupEnd, "ncclGroupEnd");
    DAFS_SYM(GetErrorString, "ncclGetErrorString"); DAFS_SYM(GetVersion, "ncclGetVersion");
#undef DAFS_SYM
    if (!ok) a->so = nullptr;
    return a;
  }();
  return api->so ? api : nullptr;
}

}  // namespace dafs

struct dafs_comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  int64_t* d_sizes = nullptr;      // [3 * (nranks + 1)]: this rank's {nnz, units, rows}, then everyone's
  int64_t* h_sizes = nullptr;      // pinned, [3 * nranks]
};

using dafs::fail;

#define DAFS_NCCL_TRY(api, expr)                                                                   \
  do {                                                                                             \
    ncclResult_t r_ = (expr);                                                                      \
    if (r_ != ncclSuccess)                                                                         \
      return ::dafs::fail(DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,           \
                          (api)->GetErrorString(r_));                                              \
  } while (0)

extern "C" int dafs_cuda_comm_unique_id(void* id128) {
  if (!id128) return fail(DAFS_CUDA_EINVAL, "id is NULL");
  const dafs::NcclApi* api = dafs::nccl_api();
  if (!api) return fail(DAFS_CUDA_ENOTSUP, "libnccl.so.2 could not be loaded");
  static_assert(sizeof(ncclUniqueId) == DAFS_CUDA_COMM_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  DAFS_NCCL_TRY(api, api->GetUniqueId(&id));
  std::memcpy(id128, &id, sizeof id);
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_comm_create(const void* id128, int rank, int nranks, dafs_comm** out) {
  if (!out) return fail(DAFS_CUDA_EINVAL, "comm is NULL");
  *out = nullptr;
  if (!id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(DAFS_CUDA_EINVAL, "bad id / rank / nranks");
  const dafs::NcclApi* api = dafs::nccl_api();
  if (!api) return fail(DAFS_CUDA_ENOTSUP, "libnccl.so.2 could not be loaded");
  dafs_comm* c = new dafs_comm();
  c->rank = rank; c->nranks = nranks;
  auto bail = [&](int rc) { dafs_cuda_comm_destroy(c); return rc; };
  if (cudaGetDevice(&c->device) != cudaSuccess) return bail(fail(DAFS_CUDA_ENODEV, "no current CUDA device"));
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof id);
  {
    ncclResult_t r = api->CommInitRank(&c->comm, nranks, id, rank);
    if (r != ncclSuccess) { c->comm = nullptr; return bail(fail(DAFS_CUDA_ECUDA, "ncclCommInitRank: %s", api->GetErrorString(r))); }
  }
  if (dafs::dev_malloc(&c->d_sizes, sizeof(int64_t) * 3 * (nranks + 1)) != cudaSuccess) return bail(fail(DAFS_CUDA_ENOMEM, "device allocation failed"));
  c->h_sizes = static_cast<int64_t*>(dafs::pinned_alloc(sizeof(int64_t) * 3 * (nranks + 1)));
  if (!c->h_sizes) return bail(fail(DAFS_CUDA_ENOMEM, "pinned allocation failed"));
  *out = c;
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_comm_destroy(dafs_comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  const dafs::NcclApi* api = dafs::nccl_api();
  if (c->comm && api) api->CommDestroy(c->comm);
  dafs::dev_free(c->d_sizes);
  if (c->h_sizes) dafs::pinned_free(c->h_sizes);
  delete c;
}

extern "C" int dafs_cuda_comm_info(const dafs_comm* c, int* rank, int* nranks, int* nccl_version) {
  if (!c) return fail(DAFS_CUDA_EINVAL, "comm is NULL");
  if (rank) *rank = c->rank;
  if (nranks) *nranks = c->nranks;
  if (nccl_version) { const dafs::NcclApi* api = dafs::nccl_api(); *nccl_version = 0; if (api) api->GetVersion(nccl_version); }
  return DAFS_CUDA_OK;
}

namespace dafs {
struct GatheredOwner {
  int device = 0;
  std::vector<int64_t> unit_off, nnz_off, rows_off;
  int64_t* d_nz_off = nullptr; int32_t* d_row_ptr = nullptr; uint32_t* d_col = nullptr; float* d_val = nullptr;
};
}  // namespace dafs

// Gather-v to `root`, exact sizes: one 24-byte all-gather of {nnz, units, rows}, then every
// non-root rank sends its four arrays and the root receives them at their offsets (one NCCL
// group: the transfers of all ranks run side by side through NVSwitch); the root's own
// fragment is a device-to-device copy.  No padding, no staging copies, nothing delivered to
// ranks that do not consume it.
extern "C" int dafs_cuda_pairhmm_plan_gather(dafs_pairhmm_plan* p, dafs_comm* c, int root, dafs_gathered_csr* out) {
  using namespace dafs;
  if (!p || !c || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_gathered_csr{};
  if (root < 0 || root >= c->nranks) return fail(DAFS_CUDA_EINVAL, "root %d outside [0, %d)", root, c->nranks);
  if (p->device != c->device) return fail(DAFS_CUDA_EINVAL, "plan and communicator live on different devices");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was handed over to a device batch");
  const NcclApi* api = nccl_api();
  if (!api) return fail(DAFS_CUDA_ENOTSUP, "libnccl.so.2 could not be loaded");
  const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
  if (rc) return rc;
  DAFS_CUDA_TRY(cudaSetDevice(c->device));
  const int R = c->nranks;
  cudaStream_t st = p->stream;
  c->h_sizes[3 * R + 0] = p->nnz; c->h_sizes[3 * R + 1] = p->npairs; c->h_sizes[3 * R + 2] = p->total_rows;
  DAFS_CUDA_TRY(cudaMemcpyAsync(c->d_sizes, c->h_sizes + 3 * R, sizeof(int64_t) * 3, cudaMemcpyHostToDevice, st));
  DAFS_NCCL_TRY(api, api->AllGather(c->d_sizes, c->d_sizes + 3, 3 * sizeof(int64_t), ncclChar, c->comm, st));
  DAFS_CUDA_TRY(cudaMemcpyAsync(c->h_sizes, c->d_sizes + 3, sizeof(int64_t) * 3 * R, cudaMemcpyDeviceToHost, st));
  DAFS_CUDA_TRY(cudaStreamSynchronize(st));
  if (c->rank != root) {
    DAFS_NCCL_TRY(api, api->GroupStart());
    DAFS_NCCL_TRY(api, api->Send(p->d_nz_off, sizeof(int64_t) * (p->npairs + 1), ncclChar, root, c->comm, st));
    if (p->total_rows) DAFS_NCCL_TRY(api, api->Send(p->d_row_ptr, sizeof(int32_t) * p->total_rows, ncclChar, root, c->comm, st));
    if (p->nnz) {
      DAFS_NCCL_TRY(api, api->Send(p->d_col, sizeof(uint32_t) * p->nnz, ncclChar, root, c->comm, st));
      DAFS_NCCL_TRY(api, api->Send(p->d_val, sizeof(float) * p->nnz, ncclChar, root, c->comm, st));
    }
    DAFS_NCCL_TRY(api, api->GroupEnd());
    DAFS_CUDA_TRY(cudaStreamSynchronize(st));
    return DAFS_CUDA_OK;
  }
  GatheredOwner* o = new GatheredOwner();
  o->device = c->device;
  o->unit_off.assign(R + 1, 0); o->nnz_off.assign(R + 1, 0); o->rows_off.assign(R + 1, 0);
  for (int r = 0; r < R; ++r) {
    o->nnz_off[r + 1] = o->nnz_off[r] + c->h_sizes[3 * r];
    o->unit_off[r + 1] = o->unit_off[r] + c->h_sizes[3 * r + 1];
    o->rows_off[r + 1] = o->rows_off[r] + c->h_sizes[3 * r + 2];
  }
  auto bail = [&](int code) { dev_free(o->d_nz_off); dev_free(o->d_row_ptr); dev_free(o->d_col); dev_free(o->d_val); delete o; return code; };
#define G_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return bail(fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_))); } while (0)
#define G_NCCL(expr) do { ncclResult_t r_ = (expr); if (r_ != ncclSuccess) return bail(fail(DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, api->GetErrorString(r_))); } while (0)
  G_TRY(dev_malloc(&o->d_nz_off, sizeof(int64_t) * (o->unit_off[R] + R)));
  G_TRY(dev_malloc(&o->d_row_ptr, sizeof(int32_t) * std::max<int64_t>(o->rows_off[R], 1)));
  G_TRY(dev_malloc(&o->d_col, sizeof(uint32_t) * std::max<int64_t>(o->nnz_off[R], 1)));
  G_TRY(dev_malloc(&o->d_val, sizeof(float) * std::max<int64_t>(o->nnz_off[R], 1)));
  G_NCCL(api->GroupStart());
  for (int r = 0; r < R; ++r) {
    const int64_t nn = c->h_sizes[3 * r], nu = c->h_sizes[3 * r + 1], nr = c->h_sizes[3 * r + 2];
    int64_t* z = o->d_nz_off + o->unit_off[r] + r;
    if (r == root) continue;
    G_NCCL(api->Recv(z, sizeof(int64_t) * (nu + 1), ncclChar, r, c->comm, st));
    if (nr) G_NCCL(api->Recv(o->d_row_ptr + o->rows_off[r], sizeof(int32_t) * nr, ncclChar, r, c->comm, st));
    if (nn) {
      G_NCCL(api->Recv(o->d_col + o->nnz_off[r], sizeof(uint32_t) * nn, ncclChar, r, c->comm, st));
      G_NCCL(api->Recv(o->d_val + o->nnz_off[r], sizeof(float) * nn, ncclChar, r, c->comm, st));
    }
  }
  G_NCCL(api->GroupEnd());
  {
    const int r = root;
    G_TRY(cudaMemcpyAsync(o->d_nz_off + o->unit_off[r] + r, p->d_nz_off, sizeof(int64_t) * (p->npairs + 1), cudaMemcpyDeviceToDevice, st));
    if (p->total_rows) G_TRY(cudaMemcpyAsync(o->d_row_ptr + o->rows_off[r], p->d_row_ptr, sizeof(int32_t) * p->total_rows, cudaMemcpyDeviceToDevice, st));
    if (p->nnz) {
      G_TRY(cudaMemcpyAsync(o->d_col + o->nnz_off[r], p->d_col, sizeof(uint32_t) * p->nnz, cudaMemcpyDeviceToDevice, st));
      G_TRY(cudaMemcpyAsync(o->d_val + o->nnz_off[r], p->d_val, sizeof(float) * p->nnz, cudaMemcpyDeviceToDevice, st));
    }
  }
  G_TRY(cudaStreamSynchronize(st));
#undef G_TRY
#undef G_NCCL
  out->nranks = R;
  out->unit_off = o->unit_off.data(); out->nnz_off = o->nnz_off.data(); out->rows_off = o->rows_off.data();
  out->d_nz_off = o->d_nz_off; out->d_row_ptr = o->d_row_ptr; out->d_col = o->d_col; out->d_val = o->d_val;
  out->owner_ = o;
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_free_gathered(dafs_gathered_csr* g) {
  if (!g || !g->owner_) return;
  dafs::GatheredOwner* o = static_cast<dafs::GatheredOwner*>(g->owner_);
  cudaSetDevice(o->device);
  dafs::dev_free(o->d_nz_off); dafs::dev_free(o->d_row_ptr); dafs::dev_free(o->d_col); dafs::dev_free(o->d_val);
  delete o;
  *g = dafs_gathered_csr{};
}

extern "C" int dafs_cuda_gathered_to_host(const dafs_gathered_csr* g, int64_t* nz_off, int32_t* row_ptr, uint32_t* col, float* val) {
  if (!g || !g->owner_) return fail(DAFS_CUDA_EINVAL, "empty gathered result");
  const int R = g->nranks;
  DAFS_CUDA_TRY(cudaSetDevice(static_cast<dafs::GatheredOwner*>(g->owner_)->device));
  if (nz_off) DAFS_CUDA_TRY(cudaMemcpy(nz_off, g->d_nz_off, sizeof(int64_t) * (g->unit_off[R] + R), cudaMemcpyDeviceToHost));
  if (row_ptr && g->rows_off[R]) DAFS_CUDA_TRY(cudaMemcpy(row_ptr, g->d_row_ptr, sizeof(int32_t) * g->rows_off[R], cudaMemcpyDeviceToHost));
  if (col && g->nnz_off[R]) DAFS_CUDA_TRY(cudaMemcpy(col, g->d_col, sizeof(uint32_t) * g->nnz_off[R], cudaMemcpyDeviceToHost));
  if (val && g->nnz_off[R]) DAFS_CUDA_TRY(cudaMemcpy(val, g->d_val, sizeof(float) * g->nnz_off[R], cudaMemcpyDeviceToHost));
  return DAFS_CUDA_OK;
}
