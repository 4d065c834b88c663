// TEST INFRASTRUCTURE — hand-written stand-in for the header the reference's cmake generates from src/config/config.hpp.in
// (CMakeLists.txt:996-1005), for the T1 oracle build (oracle/ref_t1.cpp): MPI on (the single-rank shim), no OpenMP.
#ifndef KMI_REFCFG_T1_BLISS_CONFIG_HPP
#define KMI_REFCFG_T1_BLISS_CONFIG_HPP
#define USE_MPI
#define PROJ_SRC_DIR "/root/reference"
#define PROJ_BIN_DIR "/tmp"
#define OMP_SHARE_DEFAULT
#define BLISS_UNUSED(x) do { (void)(x); } while (0)
#define BLISS_ALIGNED_ARRAY(name, count, alignsize) name alignas(alignsize) [count]
#define BLISS_ALIGNED_VAR(name, alignsize) name alignas(alignsize)
#endif
