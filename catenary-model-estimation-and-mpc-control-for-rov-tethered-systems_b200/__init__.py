"""catenary-b200: the catenary tether shape solve + RK4 theta/gamma rollout + sampling-MPC step of
eather0056/Catenary-Model-Estimation-and-MPC-Control-for-ROV-Tethered-Systems as hand-written CUDA (sm_100a)
behind the reference's Python call surface.  NumPy in / NumPy out through a C ABI (include/catenary_b200.h).
There is no CPU fallback: importing works anywhere, calling needs the built library and a CUDA device."""
from ._lib import (CatenaryLibraryError, LIB_PATH, OUT_FINAL, OUT_ORIGINAL, OUT_THETA_ALIGNED,
                   OUT_THETA_ROTATED, PROBLEM_DTYPE)
from .context import Context, default_context
from .dynamics import (EquationModel, equation_id, equation_table, extract_features, integrate_theta_gamma,
                       predict, rk4_integration)
from .geometry import (Catenary, catenary_equation, catenary_shapes, compute_catenary_3D, generate_catenary,
                       solve_catenary_params, lowest_point, rodrigues_rotation, solve_catenary, span_terms,
                       transform_catenary)
from .mpc import (circle_reference, default_mpc_params, make_params, make_problems, mpc_fleet_step,
                  mpc_sample_step, reduce_winners)
from .controller import RecedingHorizonController, mpc_receding_horizon
from .expr import (ExpressionError, ExpressionModel, compile_expression, integrate_second_order, load_equation_table,
                   load_run)
from .features import (build_clean_features, build_gamma_features, build_gamma_features_valid,
                       build_theta_features, build_theta_features_valid, compute_derivatives, extract_features16,
                       feature_bank, features_dd, gaussian_filter1d, gradient, load_and_resample_all,
                       load_trajectory_csv, preprocess_signals, resample_linear, savgol_filter, uniform_resample)

__all__ = [n for n in dir() if not n.startswith("_")]
