// compact.cu — the 24-byte upload form of SignalRef records (lsg_event24) and its widening pass.
//
// End to end a batch costs its upload: 10 M events x 32 B at the ~54 GB/s a pinned host buffer gets over PCIe is 6 ms,
// six times the kernels.  The reference has nothing to say about a wire format (its SignalRef is a Go struct of
// strings, pkg/correlation/dns.go:30-40); the interner that turns strings into ids can as well emit 24 bytes per
// event, which is what crosses PCIe.  The first device pass widens them into the 32-byte records every kernel reads
// (56 B of HBM traffic per event, hidden behind the next batch's upload).  The limits the narrow fields impose are in
// include/llmslo_gpu.h next to the struct; a record outside them is reported, never silently truncated.
#include "common.cuh"

namespace {

constexpr unsigned TS_BITS = LSG_E24_TS_BITS, IDENT_BITS = LSG_E24_IDENT_BITS, TRACE_BITS = LSG_E24_TRACE_BITS,
                   CONN_BITS = LSG_E24_CONN_BITS, PID_BITS = LSG_E24_PID_BITS;
static_assert(TS_BITS + IDENT_BITS + 4 + 4 == 64, "word a: time offset, identity row, signal, low pid bits");
static_assert(TRACE_BITS + CONN_BITS + (PID_BITS - 4) == 64, "word b: trace id, conn id, high pid bits");
constexpr uint64_t TS_MASK = ((uint64_t)1 << TS_BITS) - 1;

LSG_HD lsg_event widen_one(const lsg_event24 &c, int64_t base) {
  lsg_event e;
  const uint64_t off = c.a & TS_MASK;
  e.ts_ns = off == TS_MASK ? LSG_TS_ZERO : (int64_t)((uint64_t)base + off);
  e.value = c.value;
  e.ident = (uint32_t)((c.a >> TS_BITS) & (((uint64_t)1 << IDENT_BITS) - 1));
  const uint32_t sig = (uint32_t)((c.a >> (TS_BITS + IDENT_BITS)) & 15u);
  const uint32_t pid_lo = (uint32_t)(c.a >> 60);
  const uint32_t trace = (uint32_t)(c.b & (((uint64_t)1 << TRACE_BITS) - 1));
  e.conn = (uint32_t)((c.b >> TRACE_BITS) & (((uint64_t)1 << CONN_BITS) - 1));
  e.pid = ((uint32_t)(c.b >> (TRACE_BITS + CONN_BITS)) << 4) | pid_lo;
  e.trace_sig = (trace << 4) | sig;
  return e;
}

__global__ void __launch_bounds__(256)
k_widen_events24(const lsg_event24 *__restrict__ in, size_t n, int64_t base, lsg_event *__restrict__ out) {
  // 24-byte records: three 8-byte loads per record (a warp reads 768 contiguous bytes), one 256-bit store
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long *p = reinterpret_cast<const unsigned long long *>(in + i);
    lsg_event24 c;
    c.value = __longlong_as_double((long long)__ldg(p));
    c.a = __ldg(p + 1);
    c.b = __ldg(p + 2);
    const lsg_event e = widen_one(c, base);
    const uint64_t vb = (uint64_t)__double_as_longlong(e.value);
    U8 u;
    u.a = make_uint4((uint32_t)(uint64_t)e.ts_ns, (uint32_t)((uint64_t)e.ts_ns >> 32), (uint32_t)vb, (uint32_t)(vb >> 32));
    u.b = make_uint4(e.ident, e.pid, e.conn, e.trace_sig);
    st_v8(out + i, u);
  }
}

}  // namespace

int lsg_launch_widen_events24(lsg_ctx *ctx, const lsg_event24 *in, size_t n, int64_t base, lsg_event *out) {
  if (n == 0) return 0;
  size_t blocks = (n + 255) / 256, cap = (size_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  ProfScope ps(ctx, PK_WIDEN);
  k_widen_events24<<<(unsigned)blocks, 256, 0, ctx->stream>>>(in, n, base, out);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

extern "C" {

// Host-side packing for callers that hold 32-byte records (an interner would emit the 24-byte form directly).
// Returns the number of records that do not fit the narrow fields (their slots are written as records that match
// nothing but still count as supported or unsupported events exactly like the original), or < 0 on a bad argument.
long long lsg_pack_events24(const lsg_event *events, size_t n, int64_t base, lsg_event24 *out) {
  if (n && (!events || !out)) return LSG_ERR_INVALID;
  long long misfit = 0;
  for (size_t i = 0; i < n; ++i) {
    const lsg_event &e = events[i];
    const uint32_t trace = e.trace_sig >> 4, sig = e.trace_sig & 15u;
    uint64_t off = TS_MASK;   // the zero Time, and every time the offset cannot express: such an event matches nothing
    bool fits = true;
    if (e.ts_ns != LSG_TS_ZERO && e.ts_ns >= base && (uint64_t)e.ts_ns - (uint64_t)base < TS_MASK)
      off = (uint64_t)e.ts_ns - (uint64_t)base;
    if (e.ident >> IDENT_BITS || e.pid >> PID_BITS || e.conn >> CONN_BITS || trace >> TRACE_BITS) fits = false;
    lsg_event24 c;
    c.value = e.value;
    if (!fits) {
      // keep the signal (supported / unsupported is counted per event) and make the record unmatchable: zero Time
      misfit++;
      c.a = TS_MASK | ((uint64_t)sig << (TS_BITS + IDENT_BITS));
      c.b = 0;
    } else {
      c.a = off | ((uint64_t)e.ident << TS_BITS) | ((uint64_t)sig << (TS_BITS + IDENT_BITS)) | ((uint64_t)(e.pid & 15u) << 60);
      c.b = (uint64_t)trace | ((uint64_t)e.conn << TRACE_BITS) | ((uint64_t)(e.pid >> 4) << (TRACE_BITS + CONN_BITS));
    }
    out[i] = c;
  }
  return misfit;
}

// the inverse, for tests and for callers that want to look at what the device will see
int lsg_unpack_events24(const lsg_event24 *in, size_t n, int64_t base, lsg_event *out) {
  if (n && (!in || !out)) return LSG_ERR_INVALID;
  for (size_t i = 0; i < n; ++i) out[i] = widen_one(in[i], base);
  return 0;
}

int lsg_widen_events24_dev(lsg_ctx *ctx, const lsg_event24 *in, size_t n, int64_t base, lsg_event *out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!in || !out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  LSG_TRY(lsg_launch_widen_events24(ctx, in, n, base, out));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

}  // extern "C"
