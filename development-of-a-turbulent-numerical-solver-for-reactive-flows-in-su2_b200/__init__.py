"""B200-native reactive-flow residual hot path (drop-in behind SU2's CSolver/CNumerics API for the reactive solvers).

The product is csrc/ (hand-written FP64 CUDA for sm_100a + the C ABI of include/rx_cuda.h, built into librx_cuda.so).
This Python package is the host-side mirror of the reference's solver interface used by the tests and bench.py; the
C++ mirror compiled with SU2 lives in csrc/host/.  There is no CPU fallback: without librx_cuda.so or without a CUDA
device every entry point raises.
"""
from .rx_solver import CReactiveNSSolverGPU, RxError, load_library, build_library, LIB_PATH, source_points, residual_rms_static, libm_eval, qr_eval, merge_restart_pieces  # noqa: F401
from .geometry import read_su2, write_su2, su2_dual_mesh, su2_dual_mesh_native, load_su2_mesh, wall_distance  # noqa: F401
from .mesh import channel_mesh, channel_strip_mesh, strip_short_side_slow, partition_strips, partition_graph, partition_graph_rank, partition_mesh, partition_rank, graph_owner, strip_owner, rcm_order, renumber_points  # noqa: F401
from .synthetic import load_mechanism, default_opts, synthetic_state, synthetic_points  # noqa: F401
