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
  static const Ops iden = OpsImpl<KMI_K, 0>::make_with_dist_hash<3>();
  static const Ops cppstd = OpsImpl<KMI_K, 0>::make_with_dist_hash<4>();
  return hash == 0 ? &murmur : hash == 1 ? &farm : hash == 2 ? &farm_dbg : hash == 3 ? &iden : hash == 4 ? &cppstd : nullptr;
}
}  // namespace kmi

#ifdef KMI_MODULE
// Built as its own shared object (kmods/libkmi_k<K>.so): libkmi.so carries the common k-mer lengths and loads any other K
// on demand (kmi_create -> find_ops -> dlopen), so bliss::common::Kmer<KMER_SIZE, ...> being a template over ANY K
// (reference src/common/kmer.hpp:128-139) does not mean shipping sixty instantiations in one library.
extern "C" const kmi::Ops* kmi_ops_module(int hash) { return kmi::KMI_CAT(kmi_ops_k, KMI_K)(hash); }
extern "C" int kmi_ops_module_k(void) { return KMI_K; }
extern "C" unsigned long kmi_ops_module_abi(void) { return (unsigned long)sizeof(kmi::Ops); }
#endif
