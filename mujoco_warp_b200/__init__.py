"""mujoco_warp_b200 -- B200-native (sm_100a) implementation of MuJoCo Warp's batched physics step.

Public surface mirrors `mujoco_warp` (reference mujoco_warp/__init__.py:25-112) for the step path:
`put_model / make_data / put_data / get_data_into / reset_data / step / forward / step1 / step2`
and every stage function `f(m, d)`.
"""

from .constants import BiasType as BiasType
from .constants import BroadphaseFilter as BroadphaseFilter
from .constants import BroadphaseType as BroadphaseType
from .constants import ConeType as ConeType
from .constants import ConstraintState as ConstraintState
from .constants import ConstraintType as ConstraintType
from .constants import ContactType as ContactType
from .constants import DisableBit as DisableBit
from .constants import DynType as DynType
from .constants import EnableBit as EnableBit
from .constants import EqType as EqType
from .constants import GainType as GainType
from .constants import GeomType as GeomType
from .constants import IntegratorType as IntegratorType
from .constants import JacobianType as JacobianType
from .constants import JointType as JointType
from .constants import SolverType as SolverType
from .constants import State as State
from .constants import TrnType as TrnType
from .forward import collision as collision
from .forward import com_pos as com_pos
from .forward import crb as crb
from .forward import ctrl_noise as ctrl_noise
from .forward import euler as euler
from .forward import factor_m as factor_m
from .forward import forward as forward
from .forward import fwd_acceleration as fwd_acceleration
from .forward import fwd_actuation as fwd_actuation
from .forward import fwd_position as fwd_position
from .forward import fwd_velocity as fwd_velocity
from .forward import implicit as implicit
from .forward import kinematics as kinematics
from .forward import make_constraint as make_constraint
from .forward import mul_m as mul_m
from .forward import rungekutta4 as rungekutta4
from .forward import solve as solve
from .forward import solve_m as solve_m
from .forward import status as status
from .forward import step as step
from .forward import step1 as step1
from .forward import step2 as step2
from .forward import transmission as transmission
from .io import get_data_into as get_data_into
from .io import make_data as make_data
from .io import make_host_data as make_host_data
from .io import override_model as override_model
from .io import put_data as put_data
from .io import put_model as put_model
from .io import reset_data as reset_data
from .mjcf import load_mjcf as load_mjcf
from .mjcf import load_mjcf_string as load_mjcf_string
from .support import contact_force as contact_force
from .support import get_state as get_state
from .support import set_state as set_state
from .support import state_size as state_size
from .types import Constraint as Constraint
from .types import Contact as Contact
from .types import Data as Data
from .types import Model as Model
from .types import Option as Option
from .types import Statistic as Statistic
