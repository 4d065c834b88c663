// kmi_ops_inst.cu — compiled once per supported k-mer length: nvcc -DKMI_K=<K> ... -o kmi_ops_k<K>.o
// Exposes kmi_ops_k<K>(hash) returning the launcher table for (K, hash).
#include "kmi_ops.cuh"

#ifndef KMI_K
#error "compile with -DKMI_K=<k-mer length>"
#endif
#define KMI_CAT2(a, b) a##b
#define KMI_CAT(a, b) KMI_CAT2(a, b)

namespace kmi {
const Ops* KMI_CAT(kmi_ops_k, KMI_K)(int hash) {
  static const Ops murmur = OpsImpl<KMI_K, 0>::make();
  static const Ops farm = OpsImpl<KMI_K, 1>::make();
  static const Ops farm_dbg = OpsImpl<KMI_K, 2>::make();
  return hash == 0 ? &murmur : hash == 1 ? &farm : hash == 2 ? &farm_dbg : nullptr;
}
}  // namespace kmi
