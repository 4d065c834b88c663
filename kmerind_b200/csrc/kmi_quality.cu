// kmi_quality.cu — the PositionQualityIndex: k-mer -> (position id, quality score).
//   KmerPositionQualityTupleParser      src/io/kmer_parser.hpp:578-760     (tuple = (k-mer, (ShortSequenceKmerId, float)))
//   QualityScoreSlidingWindow           src/index/quality_score_iterator.hpp:60-140
//   Illumina18QualityScoreCodec<float>  src/index/quality_scores.hpp:60-170,529   (Phred+33: LUT[c - 33] = log2(1 - 10^(-q/10)))
//   PositionQualityIndex<MapType>       src/index/kmer_index.hpp:406
// A k-mer's score is exp2 of the sum of its K bases' log2(probability correct), or 0 when any base has a Phred score
// outside 1..93. The reference keeps a ROLLING float sum per read (subtract the leaving value, add the entering one), so
// the value of the window that starts at base i depends on the order of the i updates before it. k_fastq_quality
// reproduces exactly that: one thread per k-mer replays the read's window from its first base (or from the start of the
// N-free run when N splits sequences) up to its own start, in the reference's operation order — float add / subtract are
// IEEE operations on both sides, so the sums are bit-identical; the final exp2f is the device's (2 ulp), which is the only
// tolerance in this path (tests: rtol 1e-6 against the reference-checked oracle).
// The index itself is a composition: a POSITION-mode kmi_index whose stored value is the ORDINAL of the tuple, plus two
// side arrays (id, quality) indexed by ordinal — no kernel of the hash index changes.
#include "kmi_internal.cuh"

#include <cmath>
#include <new>

using namespace kmi;

namespace {

__constant__ float c_qlut[96];
float h_qlut[96];
bool h_qlut_ready = false;
bool d_qlut_ready[64] = {false};

void qlut_host() {
  if (h_qlut_ready) return;
  h_qlut[0] = -3.402823466e+38F;   // numeric_limits<float>::lowest(): Phred 0
  for (int q = 1; q <= 93; ++q) h_qlut[q] = (float)(std::log1p(-std::pow(10.0, -(double)q / 10.0)) / std::log(2.0));
  h_qlut[94] = h_qlut[95] = 0.0f;
  h_qlut_ready = true;
}
int qlut_device(int device) {
  qlut_host();
  if (device < 64 && d_qlut_ready[device]) return KMI_OK;
  CU(cudaMemcpyToSymbol(c_qlut, h_qlut, sizeof(h_qlut)));
  if (device < 64) d_qlut_ready[device] = true;
  return KMI_OK;
}

__device__ __forceinline__ bool q_eol(uint8_t c) { return c == '\n' || c == '\r'; }
__device__ __forceinline__ bool q_isn(uint8_t c) { return c == 'N' || c == 'n'; }
// LUT value of a quality byte; ok = "correct base" (strictly between LUT[0] and LUT[95] = 0)
__device__ __forceinline__ float q_decode(uint8_t c, bool& ok) {
  if (c < 33 || c > 128) { ok = false; return 0.0f; }
  const float v = c_qlut[c - 33];
  ok = (v > c_qlut[0]) && (v < 0.0f);
  return v;
}

// one thread per (k-mer, id) tuple of kmi_parse_fastq_dev over the same bytes
__global__ void __launch_bounds__(256)
k_fastq_quality(const uint8_t* __restrict__ bytes, u64 n, u64 file_offset, const u64* __restrict__ ids, u64 n_ids, int K, int nsplit,
                float* __restrict__ quals) {
  for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < n_ids; t += (u64)gridDim.x * blockDim.x) {
    const u64 id = ids[t];
    const u64 rec = (((id >> 16) - file_offset) & 0xFFFFFFFFFFULL);   // the record's '@' in the buffer
    const u64 kp = rec + (id & 0xFFFFull);                            // the k-mer's first base
    float out = 0.0f;
    if (rec < n && kp + (u64)K <= n) {
      u64 p = rec;
      while (p < n && !q_eol(__ldg(bytes + p))) ++p;                  // line 1
      while (p < n && q_eol(__ldg(bytes + p))) ++p;
      const u64 s0 = p;                                               // sequence line
      u64 e = kp + (u64)K;
      while (e < n && !q_eol(__ldg(bytes + e))) ++e;
      while (e < n && q_eol(__ldg(bytes + e))) ++e;                   // '+' line
      while (e < n && !q_eol(__ldg(bytes + e))) ++e;
      while (e < n && q_eol(__ldg(bytes + e))) ++e;
      const u64 q0 = e;                                               // quality line
      u64 qe = q0;
      u64 b = s0;                                                     // where the reference initialised this window's run
      if (nsplit) {
        b = kp;
        while (b > s0 && !q_isn(__ldg(bytes + b - 1))) --b;
      }
      // quality bytes available (a truncated last record may have none)
      const u64 need = (kp - s0) + (u64)K;                            // quality bytes [0, need) are read
      qe = q0 + need;
      bool have = qe <= n;
      if (have) {
        for (u64 j = q0; j < qe; ++j) if (q_eol(__ldg(bytes + j))) { have = false; break; }
      }
      if (have) {
        const uint8_t* q = bytes + q0 + (b - s0);
        float sum = 0.0f;
        int bad = 0;
        for (int j = 0; j < K; ++j) {
          bool ok;
          const float v = q_decode(__ldg(q + j), ok);
          if (ok) sum += v; else ++bad;
        }
        const u64 steps = kp - b;
        for (u64 s = 0; s < steps; ++s) {
          bool ok_o, ok_n;
          const float o = q_decode(__ldg(q + s), ok_o);
          const float v = q_decode(__ldg(q + s + K), ok_n);
          if (ok_o) sum -= o; else --bad;
          if (ok_n) sum += v; else ++bad;
        }
        out = bad > 0 ? 0.0f : exp2f(sum);
      }
    }
    quals[t] = out;
  }
}

__global__ void __launch_bounds__(256) k_iota_u64(u64* __restrict__ v, u64 n, u64 first) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) v[i] = first + i;
}
__global__ void __launch_bounds__(256)
k_gather_posqual(const u64* __restrict__ ord, u64 n, const u64* __restrict__ ids, const float* __restrict__ quals, u64* __restrict__ out_ids,
                 float* __restrict__ out_quals) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    const u64 o = ord[i];
    out_ids[i] = ids[o];
    out_quals[i] = quals[o];
  }
}

}  // namespace

extern "C" int kmi_quality_lut(float* out96) {
  if (!out96) return fail(KMI_ERR_ARG, "null argument");
  qlut_host();
  for (int i = 0; i < 96; ++i) out96[i] = h_qlut[i];
  return KMI_OK;
}

extern "C" int kmi_fastq_quality_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, const uint64_t* d_ids, size_t n_ids,
                                     float* d_quals) {
  if (!idx || (!d_bytes && n) || ((!d_ids || !d_quals) && n_ids)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  OK(qlut_device(idx->cfg.device));
  if (!n_ids) return KMI_OK;
  k_fastq_quality<<<grid_for(n_ids, 256, 148u * 32u), 256, 0, idx->stream>>>(d_bytes, (u64)n, file_offset & 0xFFFFFFFFFFULL, reinterpret_cast<const u64*>(d_ids),
                                                                          (u64)n_ids, idx->cfg.k, idx->cfg.nsplit ? 1 : 0, d_quals);
  idx->launches++;
  CU(cudaGetLastError());
  return KMI_OK;
}

// ---- the index --------------------------------------------------------------------------------------------------------------
struct kmi_posqual {
  kmi_index* base = nullptr;     // POSITION index: k-mer -> ordinals of its tuples
  DevBuf ids, quals;             // by ordinal
  u64 n = 0;                     // tuples ever appended (erased k-mers leave their rows behind)
  DevBuf text, keys, ord, out_ids, out_quals;
};

namespace {
int pq_reserve(kmi_posqual* pq, u64 extra) {
  const u64 need = pq->n + extra;
  if (need * sizeof(u64) > pq->ids.cap || need * sizeof(float) > pq->quals.cap) {
    DevBuf ni, nq;
    const u64 want = need + need / 4 + 1024;
    OK(ni.ensure(want * sizeof(u64)));
    OK(nq.ensure(want * sizeof(float)));
    if (pq->n) {
      CU(cudaMemcpyAsync(ni.p, pq->ids.p, pq->n * sizeof(u64), cudaMemcpyDeviceToDevice, pq->base->stream));
      CU(cudaMemcpyAsync(nq.p, pq->quals.p, pq->n * sizeof(float), cudaMemcpyDeviceToDevice, pq->base->stream));
      CU(cudaStreamSynchronize(pq->base->stream));
    }
    pq->ids.release(); pq->quals.release();
    pq->ids = ni; pq->quals = nq;
  }
  return KMI_OK;
}
}  // namespace

extern "C" int kmi_posqual_create(const kmi_config* cfg, kmi_posqual** out) {
  if (!cfg || !out) return fail(KMI_ERR_ARG, "null argument");
  *out = nullptr;
  if (cfg->mode != KMI_MODE_POSITION) return fail(KMI_ERR_ARG, "a PositionQualityIndex is a POSITION-mode index");
  if (cfg->nranks != 1) return fail(KMI_ERR_UNSUPPORTED, "the PositionQualityIndex is one shard");
  kmi_index* base = nullptr;
  OK(kmi_create(cfg, &base));
  kmi_posqual* pq = new (std::nothrow) kmi_posqual();
  if (!pq) { kmi_destroy(base); return fail(KMI_ERR_NOMEM, "host allocation"); }
  pq->base = base;
  *out = pq;
  return KMI_OK;
}

extern "C" void kmi_posqual_destroy(kmi_posqual* pq) {
  if (!pq) return;
  cudaSetDevice(pq->base->cfg.device);
  DevBuf* bufs[] = {&pq->ids, &pq->quals, &pq->text, &pq->keys, &pq->ord, &pq->out_ids, &pq->out_quals};
  for (DevBuf* b : bufs) b->release();
  kmi_destroy(pq->base);
  delete pq;
}

extern "C" int kmi_posqual_build_fastq_dev(kmi_posqual* pq, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!pq || (!d_bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  if (n_kmers) *n_kmers = 0;
  if (!n) return KMI_OK;
  kmi_index* idx = pq->base;
  CU(cudaSetDevice(idx->cfg.device));
  const u64 cap = (u64)n / 2 + 64;
  OK(pq_reserve(pq, cap));
  OK(pq->keys.ensure(cap * sizeof(u64) * idx->nw));
  OK(pq->ord.ensure(cap * sizeof(u64)));
  uint64_t got = 0;
  uint64_t* d_ids = pq->ids.as<uint64_t>() + pq->n;
  OK(kmi_parse_fastq_dev(idx, d_bytes, n, file_offset, pq->keys.as<uint64_t>(), d_ids, cap, &got));
  if (got) {
    OK(kmi_fastq_quality_dev(idx, d_bytes, n, file_offset, d_ids, got, pq->quals.as<float>() + pq->n));
    k_iota_u64<<<grid_for(got, 256), 256, 0, idx->stream>>>(pq->ord.as<u64>(), got, pq->n);
    idx->launches++;
    OK(kmi_insert_dev(idx, pq->keys.as<uint64_t>(), pq->ord.as<uint64_t>(), got));   // canonical keys: canonicalising again is the identity
    pq->n += got;
  }
  CU(cudaStreamSynchronize(idx->stream));
  if (n_kmers) *n_kmers = got;
  return KMI_OK;
}

extern "C" int kmi_posqual_build_fastq(kmi_posqual* pq, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!pq || (!bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(pq->base->cfg.device));
  OK(pq->text.ensure(n + 64));
  OK(copy_to_device(pq->base, pq->text.p, bytes, n));
  return kmi_posqual_build_fastq_dev(pq, pq->text.as<uint8_t>(), n, file_offset, n_kmers);
}

extern "C" int kmi_posqual_local_size(kmi_posqual* pq, uint64_t* n) {
  if (!pq) return fail(KMI_ERR_ARG, "null index");
  return kmi_local_size(pq->base, n);
}
extern "C" int kmi_posqual_count(kmi_posqual* pq, const uint64_t* keys, size_t n, uint64_t* counts) {
  if (!pq) return fail(KMI_ERR_ARG, "null index");
  return kmi_count(pq->base, keys, n, counts);
}
extern "C" int kmi_posqual_erase(kmi_posqual* pq, const uint64_t* keys, size_t n) {
  if (!pq) return fail(KMI_ERR_ARG, "null index");
  return kmi_erase(pq->base, keys, n);
}

// ordinals (host) -> ids / quals (host)
static int pq_resolve(kmi_posqual* pq, uint64_t* ids_io, float* quals, size_t m) {
  if (!m) return KMI_OK;
  kmi_index* idx = pq->base;
  OK(pq->ord.ensure(m * sizeof(u64)));
  OK(pq->out_ids.ensure(m * sizeof(u64)));
  OK(pq->out_quals.ensure(m * sizeof(float)));
  CU(cudaMemcpyAsync(pq->ord.p, ids_io, m * sizeof(u64), cudaMemcpyHostToDevice, idx->stream));
  k_gather_posqual<<<grid_for(m, 256), 256, 0, idx->stream>>>(pq->ord.as<u64>(), m, pq->ids.as<u64>(), pq->quals.as<float>(), pq->out_ids.as<u64>(),
                                                           pq->out_quals.as<float>());
  idx->launches++;
  CU(cudaMemcpyAsync(ids_io, pq->out_ids.p, m * sizeof(u64), cudaMemcpyDeviceToHost, idx->stream));
  if (quals) CU(cudaMemcpyAsync(quals, pq->out_quals.p, m * sizeof(float), cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  CU(cudaGetLastError());
  return KMI_OK;
}

extern "C" int kmi_posqual_find(kmi_posqual* pq, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* ids, float* quals, size_t cap) {
  if (!pq || !offsets) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(pq->base->cfg.device));
  OK(kmi_find_positions(pq->base, keys, n, offsets, ids, cap));
  if (!ids) return KMI_OK;
  return pq_resolve(pq, ids, quals, (size_t)offsets[n]);
}

extern "C" int kmi_posqual_dump(kmi_posqual* pq, uint64_t* keys, uint64_t* ids, float* quals, size_t cap, size_t* n_out) {
  if (!pq || !n_out || !ids) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(pq->base->cfg.device));
  OK(kmi_dump(pq->base, keys, ids, cap, n_out));
  return pq_resolve(pq, ids, quals, *n_out);
}

extern "C" int kmi_posqual_launch_count(kmi_posqual* pq, uint64_t* n) {
  if (!pq || !n) return fail(KMI_ERR_ARG, "null argument");
  *n = pq->base->launches;
  return KMI_OK;
}
