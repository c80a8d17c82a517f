// TEST INFRASTRUCTURE (oracle/): hand-written MFEM configuration header used only to
// compile the UNMODIFIED reference sources under /root/reference into oracle/_ref/.
// It replaces the cmake-generated `_config.hpp` (third_party/mfem/config/config.hpp:20-24
// honours MFEM_CONFIG_FILE) with the values mimi's own build would produce
// (/root/reference/setup.py:74-81: serial, double precision; SuiteSparse is absent here and
// therefore OFF -- mfem::UMFPackSolver is never reached by the hot path).
#ifndef MIMI_B200_ORACLE_REF_MFEM_CONFIG_HPP
#define MIMI_B200_ORACLE_REF_MFEM_CONFIG_HPP
#define MFEM_VERSION 40701
#define MFEM_VERSION_STRING "4.7.1"
#define MFEM_VERSION_TYPE ((MFEM_VERSION) % 2)
#define MFEM_VERSION_TYPE_RELEASE 0
#define MFEM_VERSION_TYPE_DEVELOPMENT 1
#define MFEM_VERSION_MAJOR ((MFEM_VERSION) / 10000)
#define MFEM_VERSION_MINOR (((MFEM_VERSION) / 100) % 100)
#define MFEM_VERSION_PATCH ((MFEM_VERSION) % 100)
#define MFEM_SOURCE_DIR "/root/reference/third_party/mfem"
#define MFEM_INSTALL_DIR "/nonexistent"
#define MFEM_GIT_STRING "0952809"
#define MFEM_USE_DOUBLE
#define MFEM_USE_MEMALLOC
#define MFEM_TIMER_TYPE 2
#endif
