// Hand-written stand-in for the header the reference's cmake would generate from
// src/config/bliss-config.hpp.in (CMakeLists.txt:996-1005). No MPI, no OpenMP.
#ifndef KMI_REFCFG_BLISS_CONFIG_HPP
#define KMI_REFCFG_BLISS_CONFIG_HPP
#define PROJ_SRC_DIR "/root/reference"
#endif
