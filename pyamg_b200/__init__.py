"""pyamg_b200 -- B200-native (sm_100a) drop-in for PyAMG's solve-phase V-cycle hot path:
CSR/BSR SpMV, restriction/prolongation, Jacobi / multicolour Gauss-Seidel / BSR Gauss-Seidel
smoothers and the Galerkin R*A*P product, behind the reference's multilevel_solver.solve() /
aspreconditioner() API and amg_core signatures.  All arithmetic runs in libb200amg.so
(hand-written CUDA); there is no CPU fallback."""
from . import _lib
from .sparse import DevMat, matmat, transpose, galerkin
from .multilevel import multilevel_solver, coarse_grid_solver
from .smoothing import change_smoothers, rho_D_inv_A
from .aggregation import smoothed_aggregation_solver
from .classical import ruge_stuben_solver
from .hierarchy import import_hierarchy
from . import amg_core, relaxation, gallery, linalg, krylov, partition, distributed, classical

__all__ = ['DevMat', 'matmat', 'transpose', 'galerkin', 'multilevel_solver', 'coarse_grid_solver',
           'change_smoothers', 'rho_D_inv_A', 'smoothed_aggregation_solver', 'ruge_stuben_solver', 'import_hierarchy',
           'amg_core', 'relaxation', 'gallery', 'linalg', 'krylov', 'classical']
