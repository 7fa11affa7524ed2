// tb_noalloc.cu -- tb_gopher_repetition_noalloc(): the GopherRepetitionFilter pass for a small batch
// with every byte of working memory taken from one fixed array that is part of the library image.
//
// Who needs this.  Some embedders are not allowed to enter the CUDA allocator from the thread that
// calls us: a worker whose device memory is handed out, once, by a pool owned by another component
// (a serving process with a fixed per-tenant quota, a host runtime that pre-partitions the GPU), or
// code on a latency-critical path where a cudaMalloc -- which may synchronise the device and take a
// driver-wide lock -- is simply forbidden.  The regular entry points (tb_pipeline_run_batch*) grow
// their buffers with cudaMalloc / cudaMallocHost on demand; this one never does.  The caller brings
// the inputs and the outputs as device pointers; all intermediates (document ranges, word and line
// hashes, hash-table work areas, the Unicode tables) live in `tb_noalloc_scratch` below.
//
// What it costs.  The array is TB_NOALLOC_SCRATCH_BYTES (24 MiB) of device memory on every GPU that
// loads the library, for the life of the process.  That bounds the batch: at most
// TB_NOALLOC_MAX_DOCS documents and TB_NOALLOC_MAX_BYTES bytes of text per call; larger requests are
// refused with TB_ERR_INVALID_ARGUMENT (split the batch, or use the allocating entry points).
//
// Rules of use.
//  * The first call on a device resolves the array's device address (the name of a __device__ array
//    is only a host-side handle; kernels must be given the address cudaGetSymbolAddress returns),
//    uploads the Unicode tables into it with cudaMemcpy and opts one kernel into its shared-memory
//    size.  That first call blocks; make it during start-up, before any stream capture is begun.
//  * Every call uses the same array and leaves its contents behind.  The few words that must start
//    at zero (tier counters and work cursors) are cleared on the caller's stream at the start of
//    each call; nothing else is assumed clean.
//  * Two calls whose GPU work overlaps (different streams, or different host threads without an
//    ordering between their streams) race on the array and produce garbage.  Serialise them.
//
// It has no counterpart in the reference and it changes none of the other code paths: it only
// launches the existing kernels through their launchers with pointers into the static array.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <string>

#include "tb_hostutil.h"
#include "tb_kernels.cuh"
#include "unicode_tables.inc"

#define TB_NOALLOC_SCRATCH_BYTES (24u << 20)

// The working memory.  In host code this symbol is NOT a device pointer; see setup_device().
static __device__ __align__(256) unsigned char tb_noalloc_scratch[TB_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
struct DeviceState {
  bool ready = false;
  unsigned char* base = nullptr;  // device address of tb_noalloc_scratch on this device
  UcTables uc{};
  size_t tables_end = 0;          // first free byte after the tables
  int sm_count = 148;
};
DeviceState g_dev[kMaxDevices];
std::mutex g_mu;

// the message goes to the caller's buffer (this entry point keeps no state of its own besides the scratch)
struct Err { char* buf; size_t len; };
thread_local Err g_errbuf{nullptr, 0};
int fail(int code, const std::string& msg) {
  if (g_errbuf.buf && g_errbuf.len) { size_t n = msg.size() < g_errbuf.len - 1 ? msg.size() : g_errbuf.len - 1; memcpy(g_errbuf.buf, msg.data(), n); g_errbuf.buf[n] = 0; }
  return code;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Once per device, with that device current: resolve the address, place and upload the tables.
int setup_device(int dev, DeviceState** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceState& st = g_dev[dev];
  if (!st.ready) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, tb_noalloc_scratch);
    if (e != cudaSuccess || !p) return fail(TB_ERR_CUDA, std::string("cudaGetSymbolAddress(tb_noalloc_scratch): ") + cudaGetErrorString(e));
    st.base = (unsigned char*)p;
    size_t off = 0;
    auto place = [&](const void* src, size_t bytes) -> unsigned char* {
      unsigned char* d = st.base + off;
      off = align_up(off + bytes, 256);
      if (cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) return nullptr;
      return d;
    };
    uint64_t pw[TB_POWTAB_ALL];
    tb_fill_powtab(pw, TB_B1_DEFAULT, TB_B2_DEFAULT);  // (a fixed pair: this entry point keeps no per-process state beyond the scratch)
    unsigned char* s1 = place(tb_uc_stage1, sizeof tb_uc_stage1);
    unsigned char* s2 = place(tb_uc_stage2, sizeof tb_uc_stage2);
    unsigned char* lo = place(tb_uc_lower, sizeof tb_uc_lower);
    unsigned char* ca = place(tb_uc_cased, sizeof tb_uc_cased);
    unsigned char* ci = place(tb_uc_caseign, sizeof tb_uc_caseign);
    unsigned char* pp = place(pw, sizeof pw);
    if (!s1 || !s2 || !lo || !ca || !ci || !pp) { cudaGetLastError(); return fail(TB_ERR_CUDA, "uploading the Unicode tables into the static scratch failed"); }
    st.uc = UcTables{(const uint16_t*)s1, (const uint16_t*)s2, (const uint32_t*)lo, (const uint32_t*)ca, (const uint32_t*)ci,
                     (const uint64_t*)pp, TB_UC_NLOWER, TB_UC_NCASED, TB_UC_NCASEIGN, {0}};
    tb_compute_wslead(tb_uc_stage1, tb_uc_stage2, st.uc.wslead);
    st.tables_end = off;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) st.sm_count = prop.multiProcessorCount;
    if (tb_kernels_init() != 0) { cudaGetLastError(); return fail(TB_ERR_CUDA, "cudaFuncSetAttribute failed (not an sm_100-class device?)"); }
    // The table uploads come from pageable memory: cudaMemcpy may return once the data is staged, with the DMA still
    // ordered on the NULL stream only.  The caller's stream may be a non-blocking one, so wait for the device here
    // (the first call is documented as blocking) before any kernel can read the tables.
    if (cudaDeviceSynchronize() != cudaSuccess) { cudaGetLastError(); return fail(TB_ERR_CUDA, "cudaDeviceSynchronize after the table upload failed"); }
    st.ready = true;
  }
  *out = &st;
  return TB_OK;
}

// Document ranges from the caller's offsets, which are device memory this entry point cannot inspect on the host
// (it must not synchronise).  The scratch is carved for n_bytes of text, so an offset outside [0, n_bytes] or a
// decreasing pair would index outside the static array: such a document is not processed.  It gets an empty range,
// keep = 0 and statistics that are all zero except flags = TB_DOCFLAG_BAD_OFFSETS.  A document [a, b) is accepted
// when hi <= a <= b <= n_bytes, hi being the end of the last accepted document before it (0 at the start): accepted
// documents are then disjoint and in order, which is what the per-document scratch layout relies on.  One CTA
// (n <= 2048); the walk over the offsets is sequential and short.
__global__ void k_noalloc_ranges(const int64_t* off, int64_t n, int64_t n_bytes, int64_t* rs, int64_t* re, uint8_t* keep,
                                 int8_t* failed, tb_gr_stats* stats) {
  __shared__ int64_t s_pos[TB_NOALLOC_MAX_DOCS];  // rejected document d: where its empty range goes (-1: accepted)
  if (threadIdx.x == 0) {
    int64_t hi = 0;
    for (int64_t d = 0; d < n; d++) {
      const int64_t a = off[d], b = off[d + 1];
      if (hi <= a && a <= b && b <= n_bytes) { s_pos[d] = -1; hi = b; }
      else s_pos[d] = hi;
    }
  }
  __syncthreads();
  for (int64_t d = threadIdx.x; d < n; d += blockDim.x) {
    if (s_pos[d] < 0) { rs[d] = off[d]; re[d] = off[d + 1]; keep[d] = 1; failed[d] = -1; }
    else {
      rs[d] = s_pos[d]; re[d] = s_pos[d]; keep[d] = 0; failed[d] = 0;
      tb_gr_stats z;
      memset(&z, 0, sizeof z);
      z.flags = TB_DOCFLAG_BAD_OFFSETS;
      stats[d] = z;
    }
  }
}

}  // namespace

extern "C" {

int tb_gopher_repetition_noalloc(const uint8_t* utf8_dev, const int64_t* offsets_dev, int64_t n_docs, int64_t n_bytes,
                                 const tb_gr_params* params, tb_gr_stats* stats_dev, uint8_t* keep_dev, void* cuda_stream,
                                 char* err, size_t err_len) {
  g_errbuf = Err{err, err_len};
  if (err && err_len) err[0] = 0;
  if (n_docs < 0 || n_bytes < 0 || !params || (n_docs > 0 && (!offsets_dev || !stats_dev || !keep_dev)) || (n_bytes > 0 && !utf8_dev))
    return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: NULL or negative argument");
  if (n_docs > TB_NOALLOC_MAX_DOCS || n_bytes > TB_NOALLOC_MAX_BYTES)
    return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: batch of " + std::to_string(n_docs) + " documents / " +
                                             std::to_string(n_bytes) + " bytes exceeds the fixed scratch (at most " +
                                             std::to_string(TB_NOALLOC_MAX_DOCS) + " documents and " + std::to_string(TB_NOALLOC_MAX_BYTES) +
                                             " bytes per call); split the batch or use tb_pipeline_run_batch");
  if (params->n_top < 0 || params->n_top > TB_MAX_NGRAMS || params->n_dup < 0 || params->n_dup > TB_MAX_NGRAMS)
    return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: at most " + std::to_string(TB_MAX_NGRAMS) + " entries per n-gram list");
  for (int k = 0; k < params->n_top; k++) if (params->top_n[k] == 0 || params->top_n[k] > TB_MAX_NGRAM_N) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: n-gram order out of range");
  for (int k = 0; k < params->n_dup; k++) if (params->dup_n[k] == 0 || params->dup_n[k] > TB_MAX_NGRAM_N) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: n-gram order out of range");
  if (n_docs == 0) return TB_OK;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) { cudaGetLastError(); return fail(TB_ERR_CUDA, "no current CUDA device"); }
  DeviceState* st = nullptr;
  int rc = setup_device(dev, &st);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)cuda_stream;

  // carve the per-call arrays out of the static array (nothing here allocates)
  const size_t n1 = (size_t)n_docs + 1, wunits = tb_word_units((size_t)n_bytes, (size_t)n_docs), units = tb_line_units((size_t)n_bytes, (size_t)n_docs);
  size_t off = st->tables_end;
  auto carve = [&](size_t bytes) -> unsigned char* { unsigned char* p = st->base + off; off = align_up(off + bytes, 256); return p; };
  int64_t* rs = (int64_t*)carve(n1 * 8); int64_t* re = (int64_t*)carve(n1 * 8);
  int64_t* ts = (int64_t*)carve(n1 * 8); int64_t* te = (int64_t*)carve(n1 * 8);
  int8_t* failed = (int8_t*)carve(n1);
  GrScratch g;
  g.wh = (uint64_t*)carve(wunits * 8); g.wl = (uint32_t*)carve(wunits * 4); g.ws = (uint64_t*)carve(wunits * 8);
  g.lh = (uint64_t*)carve(units * 8); g.ll = (uint32_t*)carve(units * 4); g.lg = (uint8_t*)carve(units);
  g.doc = (GrDoc*)carve(n1 * sizeof(GrDoc));
  g.work = (uint32_t*)carve((wunits * TB_REP_K + n1 * TB_REP_DOC_EXTRA) * 4);
  g.lists = (uint32_t*)carve(TB_REP_NT * n1 * 4);
  g.counters = (uint32_t*)carve(256);  // [TB_GR_COUNTERS, +16): the counts of the work order
  uint32_t* order = (uint32_t*)carve(n1 * 4);
  uint8_t* sizecls = (uint8_t*)carve(n1);
  g.n = n_docs;
  if (off > TB_NOALLOC_SCRATCH_BYTES) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gopher_repetition_noalloc: internal scratch budget exceeded");

  GrCfg cfg{};
  cfg.has = params->has; cfg.n_top = params->n_top; cfg.n_dup = params->n_dup;
  cfg.dup_line_frac = params->dup_line_frac; cfg.dup_para_frac = params->dup_para_frac;
  cfg.dup_line_char_frac = params->dup_line_char_frac; cfg.dup_para_char_frac = params->dup_para_char_frac;
  for (int k = 0; k < TB_MAX_NGRAMS; k++) { cfg.top_n[k] = params->top_n[k]; cfg.top_frac[k] = params->top_frac[k]; cfg.dup_n[k] = params->dup_n[k]; cfg.dup_frac[k] = params->dup_frac[k]; }
  tb_grcfg_finish(cfg);

  // state that must start at zero: tier list lengths and work cursors (the array keeps the last call's contents)
  if (cudaMemsetAsync(g.counters, 0, 256, stream) != cudaSuccess ||
      cudaMemsetAsync(stats_dev, 0, (size_t)n_docs * sizeof(tb_gr_stats), stream) != cudaSuccess) {  // (pass 2 writes the non-zero fields only)
    cudaGetLastError(); return fail(TB_ERR_CUDA, "cudaMemsetAsync failed");
  }
  // keep_dev doubles as the kernels' "alive" flag: 1 for every document whose offsets are usable
  k_noalloc_ranges<<<1, 1024, 0, stream>>>(offsets_dev, n_docs, n_bytes, rs, re, keep_dev, failed, stats_dev);
  tb_launch_order(stream, rs, re, n_docs, order, g.counters + TB_GR_COUNTERS, sizecls);
  KBatch b{utf8_dev, rs, re, keep_dev, failed, n_docs, st->uc, 0, n_bytes, 0u, order, g.counters + TB_GR_COUNTERS, sizecls, 0u};
  tb_launch_trim(stream, b, ts, te);
  tb_launch_gr_scan(stream, b, ts, te, g, st->sm_count);
  for (int tier = 0; tier < TB_REP_NT; tier++) tb_launch_gr_rep(stream, b, ts, te, g, cfg, 0, stats_dev, st->sm_count, tier);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(TB_ERR_CUDA, std::string("tb_gopher_repetition_noalloc: launch failed: ") + cudaGetErrorString(e));
  return TB_OK;
}

}  // extern "C"
