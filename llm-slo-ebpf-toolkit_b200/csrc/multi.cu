// multi.cu — the path on several GPUs of one box behind the C ABI (SURVEY.md §8b "multi-GPU variants take per-rank
// arrays", §8e): one lsg_ctx per device inside one process, the batch sharded by the caller (join groups never split:
// shard by node / pod identity), every device runs the whole single-GPU pipeline on its shard, and the only exchange is
// one ncclAllReduce(ncclUint64, ncclSum) of the counter block (debug counters, confusion matrix, accuracy numerators).
// Ratios are formed from the reduced integers by the caller, so the result does not depend on the reduction order.
//
// NCCL is not linked: the library is looked up at run time (dlopen of libnccl.so.2) the first time lsg_multi_create
// is called, so a single-GPU user needs no NCCL at all and a process that already carries its own copy (PyTorch bundles
// one) is not disturbed.  The function table below is the whole of NCCL this file uses.
#include <dlfcn.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdlib>
#include <mutex>
#include <new>

#include "common.cuh"

void *lsg_internal_slot_counters_dev(lsg_ctx *ctx, int slot);   // api.cu: the device counter block of a pipeline slot

namespace {

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int *) = nullptr;
};
std::mutex g_nccl_mu;
NcclApi g_nccl;
std::string g_nccl_err;

bool load_nccl() {
  std::lock_guard<std::mutex> lock(g_nccl_mu);
  if (g_nccl.handle) return true;
  const char *names[] = {getenv("LSG_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) {
    if (!n || !*n) continue;
    h = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    if (h) break;
  }
  if (!h) {
    const char *e = dlerror();
    g_nccl_err = std::string("NCCL not found (libnccl.so.2; set LSG_NCCL_LIB): ") + (e ? e : "");
    return false;
  }
  NcclApi a;
  a.handle = h;
  a.CommInitAll = (decltype(a.CommInitAll))dlsym(h, "ncclCommInitAll");
  a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
  a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce");
  a.GroupStart = (decltype(a.GroupStart))dlsym(h, "ncclGroupStart");
  a.GroupEnd = (decltype(a.GroupEnd))dlsym(h, "ncclGroupEnd");
  a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
  a.GetVersion = (decltype(a.GetVersion))dlsym(h, "ncclGetVersion");
  if (!a.CommInitAll || !a.CommDestroy || !a.AllReduce || !a.GroupStart || !a.GroupEnd || !a.GetErrorString) {
    g_nccl_err = "the NCCL library lacks one of the six entry points this path uses";
    dlclose(h);
    return false;
  }
  g_nccl = a;
  return true;
}

}  // namespace

struct lsg_multi {
  int n = 0;
  std::vector<int> devices;
  std::vector<lsg_ctx *> ctx;
  std::vector<ncclComm_t> comm;
  std::vector<unsigned long long *> reduced_dev;   // per rank: the allreduced counter block
  lsg_counters *reduced_host = nullptr;            // pinned, rank 0's copy
  std::string err;
};

namespace {
int mfail(lsg_multi *m, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (m) m->err = buf;
  return code;
}
}  // namespace

extern "C" {

int lsg_multi_create(const int *devices, int n_devices, lsg_multi **out) {
  if (!out) return LSG_ERR_INVALID;
  *out = nullptr;
  if (!devices || n_devices < 1 || n_devices > 64) return LSG_ERR_INVALID;
  for (int i = 0; i < n_devices; ++i)
    for (int j = 0; j < i; ++j)
      if (devices[i] == devices[j]) return LSG_ERR_INVALID;   // one rank per device (NCCL requires it too)
  lsg_multi *m = new (std::nothrow) lsg_multi();
  if (!m) return LSG_ERR_NOMEM;
  m->n = n_devices;
  m->devices.assign(devices, devices + n_devices);
  m->ctx.assign(n_devices, nullptr);
  m->comm.assign(n_devices, nullptr);
  m->reduced_dev.assign(n_devices, nullptr);
  *out = m;   // from here on the caller can read lsg_multi_last_error and must call lsg_multi_destroy
  for (int r = 0; r < n_devices; ++r) {
    const int rc = lsg_create(devices[r], nullptr, &m->ctx[r]);
    if (rc) return mfail(m, rc, "lsg_create failed on device %d (rank %d)", devices[r], r);
    if (cudaSetDevice(devices[r]) != cudaSuccess || cudaMalloc(&m->reduced_dev[r], sizeof(lsg_counters)) != cudaSuccess) {
      cudaGetLastError();
      return mfail(m, LSG_ERR_NOMEM, "device allocation failed on device %d", devices[r]);
    }
  }
  if (cudaMallocHost(&m->reduced_host, sizeof(lsg_counters)) != cudaSuccess) {
    cudaGetLastError();
    return mfail(m, LSG_ERR_NOMEM, "pinned allocation failed");
  }
  if (!load_nccl()) return mfail(m, LSG_ERR_CUDA, "%s", g_nccl_err.c_str());
  const ncclResult_t nr = g_nccl.CommInitAll(m->comm.data(), n_devices, devices);
  if (nr != ncclSuccess) {
    for (auto &c : m->comm) c = nullptr;
    return mfail(m, LSG_ERR_CUDA, "ncclCommInitAll over %d devices failed: %s", n_devices, g_nccl.GetErrorString(nr));
  }
  return 0;
}

void lsg_multi_destroy(lsg_multi *m) {
  if (!m) return;
  for (int r = 0; r < m->n; ++r) {
    if (m->ctx[r]) {
      cudaSetDevice(m->devices[r]);
      cudaStreamSynchronize(m->ctx[r]->stream);
    }
    if (m->comm[r] && g_nccl.CommDestroy) g_nccl.CommDestroy(m->comm[r]);
    if (m->reduced_dev[r]) { cudaSetDevice(m->devices[r]); cudaFree(m->reduced_dev[r]); }
    if (m->ctx[r]) lsg_destroy(m->ctx[r]);
  }
  if (m->reduced_host) cudaFreeHost(m->reduced_host);
  delete m;
}

const char *lsg_multi_last_error(const lsg_multi *m) { return m ? m->err.c_str() : "null handle"; }
int lsg_multi_size(const lsg_multi *m) { return m ? m->n : 0; }
lsg_ctx *lsg_multi_ctx(lsg_multi *m, int rank) { return m && rank >= 0 && rank < m->n ? m->ctx[rank] : nullptr; }

int lsg_multi_nccl_version(void) {
  int v = 0;
  if (load_nccl() && g_nccl.GetVersion) g_nccl.GetVersion(&v);
  return v;
}

int lsg_multi_pipeline(lsg_multi *m, const lsg_pipeline_args *args, lsg_counters *counters_out, lsg_counters *reduced_out) {
  if (!m || !args) return LSG_ERR_INVALID;
  for (int r = 0; r < m->n; ++r)
    if (!m->ctx[r] || !m->comm[r]) return mfail(m, LSG_ERR_INVALID, "the handle was not created completely: %s", m->err.c_str());
  // every rank's upload, kernels and download are enqueued before anything waits: the devices work at the same time
  for (int r = 0; r < m->n; ++r) {
    const int rc = lsg_pipeline_submit(m->ctx[r], &args[r], 0);
    if (rc) {
      for (int q = 0; q < r; ++q) lsg_pipeline_wait(m->ctx[q], 0, nullptr);
      return mfail(m, rc, "rank %d: %s", r, lsg_last_error(m->ctx[r]));
    }
  }
  // the path's one exchange, on every rank's own stream behind its kernels
  ncclResult_t nr = g_nccl.GroupStart();
  for (int r = 0; r < m->n && nr == ncclSuccess; ++r)
    nr = g_nccl.AllReduce(lsg_internal_slot_counters_dev(m->ctx[r], 0), m->reduced_dev[r], LSG_COUNTER_WORDS, ncclUint64, ncclSum,
                          m->comm[r], m->ctx[r]->stream);
  const ncclResult_t ne = g_nccl.GroupEnd();
  if (nr == ncclSuccess) nr = ne;
  int first_rc = 0;
  if (nr != ncclSuccess) first_rc = mfail(m, LSG_ERR_CUDA, "ncclAllReduce of the counter block failed: %s", g_nccl.GetErrorString(nr));
  if (!first_rc) {
    cudaSetDevice(m->devices[0]);
    if (cudaMemcpyAsync(m->reduced_host, m->reduced_dev[0], sizeof(lsg_counters), cudaMemcpyDeviceToHost, m->ctx[0]->stream) != cudaSuccess)
      first_rc = mfail(m, LSG_ERR_CUDA, "copy of the reduced counters failed");
  }
  std::vector<lsg_counters> own((size_t)m->n);
  for (int r = 0; r < m->n; ++r) {
    const int rc = lsg_pipeline_wait(m->ctx[r], 0, &own[r]);
    if (rc && !first_rc) first_rc = mfail(m, rc, "rank %d: %s", r, lsg_last_error(m->ctx[r]));
    if (counters_out) counters_out[r] = own[r];
  }
  for (int r = 0; r < m->n; ++r) {   // the allreduce (and rank 0's copy) sit behind the kernels on each rank's stream
    cudaSetDevice(m->devices[r]);
    if (cudaStreamSynchronize(m->ctx[r]->stream) != cudaSuccess && !first_rc)
      first_rc = mfail(m, LSG_ERR_CUDA, "rank %d: stream synchronisation failed: %s", r, cudaGetErrorString(cudaGetLastError()));
  }
  if (first_rc) return first_rc;
  if (reduced_out) {
    lsg_counters red = *m->reduced_host;
    // Unmatched and UnsupportedType count pairs, and a sum over shards only sees the pairs inside each shard.  The
    // reference counts every pair of the batch (correlator.go:73-82 runs over all signals for every span): the matched
    // pairs are the shards' own (groups are never split), every other supported pair of the whole batch is unmatched.
    unsigned long long matched = 0;
    for (int r = 0; r < m->n; ++r)
      matched += (own[r].events - own[r].unsupported_events) * own[r].spans - own[r].debug.unmatched;
    red.debug.unsupported_type = red.unsupported_events * red.spans;
    red.debug.unmatched = (red.events - red.unsupported_events) * red.spans - matched;
    *reduced_out = red;
  }
  return 0;
}

}  // extern "C"
