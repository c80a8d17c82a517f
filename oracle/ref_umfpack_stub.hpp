// TEST INFRASTRUCTURE (oracle/). Forced-include used only when compiling the unmodified reference
// sources into oracle/_ref/: SuiteSparse is not installed in this image, so MFEM is configured
// without it and `mfem::UMFPackSolver` (a data member at
// /root/reference/src/mimi/integrators/nonlinear_solid.hpp:21 that the hot path never calls) has
// no definition. This gives it one that refuses to be used.
#ifndef MIMI_B200_ORACLE_REF_UMFPACK_STUB_HPP
#define MIMI_B200_ORACLE_REF_UMFPACK_STUB_HPP
#include <mfem.hpp>
#ifndef MFEM_USE_SUITESPARSE
namespace mfem {
class UMFPackSolver : public Solver {
public:
  double Control[32];
  double Info[96];
  UMFPackSolver() {}
  void SetOperator(const Operator&) override { mfem_error("UMFPackSolver stub: SuiteSparse absent"); }
  void Mult(const Vector&, Vector&) const override { mfem_error("UMFPackSolver stub: SuiteSparse absent"); }
  void SetPrintLevel(int) {}
};
} // namespace mfem
#endif
#endif
