"""Agents of the hot path with the reference's interface (`mantrap/agents`): a discrete-time agent holds its state
`(px, py, vx, vy, time)` and history on the host (scene description) and evaluates its dynamics -- single steps,
batched trajectory unrolling / rolling, `dx_du` -- through the CUDA ego kernels of the C-ABI library.

Mirrors (paths relative to the reference root):
  DTAgent                  mantrap/agents/base/discrete.py:15-575
  LinearDTAgent            mantrap/agents/base/linear.py:13-236
  IntegratorDTAgent        mantrap/agents/integrator_single.py:9-107     v' = u,          p' = p + u dt
  DoubleIntegratorDTAgent  mantrap/agents/integrator_double.py:9-121     v' = v + u dt,   p' = p + v dt (OLD v)
Out of scope here: go_to_point / p-ahead controller, expand_trajectory, reachability circle.

Tensors handed in may live on the CPU (as in the reference) or on the engine's device; results come back on the
device of the input, connected to it in the autograd graph (`EgoUnroll`).
"""
import math
import random
import string
import typing

import numpy as np
import torch

from . import constants, runtime
from .autograd import EgoUnroll


class DTAgent:
    ego_kind: str = None          # "single" | "double": key of the CUDA dynamics

    def __init__(self, position: torch.Tensor, velocity: torch.Tensor = torch.zeros(2), time: float = 0,
                 history: torch.Tensor = None, is_robot: bool = False, color: np.ndarray = None,
                 identifier: str = None, **agent_params):
        position, velocity = torch.as_tensor(position).float().cpu(), torch.as_tensor(velocity).float().cpu()
        assert position.shape == (2,) and velocity.shape == (2,), "position and velocity must be 2D vectors"
        assert time >= 0
        self._state = torch.cat((position, velocity, torch.ones(1) * time))
        row = self._state.unsqueeze(0)
        if history is not None:
            history = torch.as_tensor(history).float().cpu()
            assert history.dim() == 2 and history.shape[1] == 5
            self._history = history if torch.allclose(history[-1], self._state) else torch.cat((history, row))
        else:
            self._history = row
        self._is_robot = is_robot
        self._color = np.zeros(3) if is_robot else (color if color is not None else np.random.rand(3))
        self._params = agent_params
        self._id = identifier if identifier is not None else "".join(random.choice(string.ascii_lowercase) for _ in range(3))

    # ---- dynamics ----------------------------------------------------------------------------------------------
    def unroll_trajectory(self, controls: torch.Tensor, dt: float) -> torch.Tensor:
        """controls [T,2] -> trajectory [T+1,5] starting at the current state (linear.py:68-105)."""
        assert controls.dim() == 2 and controls.shape[1] == 2
        return self._unroll(self.state_with_time, controls, dt)

    def _unroll(self, state: torch.Tensor, controls: torch.Tensor, dt: float) -> torch.Tensor:
        engine = runtime.default_engine()
        u = runtime.to_device(controls.float(), engine)[None]
        x0 = runtime.to_device(state.detach().float(), engine)[None]
        trajectory = EgoUnroll.apply(u, x0, engine, self.ego_kind, float(dt))[0]
        return trajectory.to(controls.device)

    def roll_trajectory(self, trajectory: torch.Tensor, dt: float) -> torch.Tensor:
        """trajectory [T+1,5] -> controls [T,2] by inverse dynamics (linear.py:107-130); not differentiable here."""
        assert trajectory.dim() == 2 and trajectory.shape[1] == 5
        engine = runtime.default_engine()
        controls = engine.ego_roll(self.ego_kind, float(dt), runtime.to_device(trajectory.detach(), engine)[None])[0]
        return controls.to(trajectory.device)

    def dx_du(self, controls: torch.Tensor, dt: float) -> torch.Tensor:
        """[5(T+1), 2T] Jacobian of the unrolled trajectory w.r.t. the flattened controls (linear.py:169-196)."""
        engine = runtime.default_engine()
        return engine.ego_dx_du(self.ego_kind, float(dt), int(controls.shape[0])).to(controls.device)

    def dynamics(self, state: torch.Tensor, action: torch.Tensor, dt: float) -> torch.Tensor:
        """One step x_{k+1} = f(x_k, u_k) for a 5-vector state with time (discrete.py:92-111)."""
        assert state.shape == (5,) and action.shape == (2,)
        return self._unroll(state, action.float().view(1, 2), dt)[1]

    def inverse_dynamics(self, state: torch.Tensor, state_previous: torch.Tensor, dt: float) -> torch.Tensor:
        """Control that moves `state_previous` to `state` (discrete.py:117-135); states are 4- or 5-vectors."""
        pair = torch.zeros((2, 5))
        pair[0, :state_previous.numel()] = state_previous.detach().float().cpu()
        pair[1, :state.numel()] = state.detach().float().cpu()
        return self.roll_trajectory(pair, dt)[0]

    # ---- state updates -----------------------------------------------------------------------------------------
    def update(self, action: torch.Tensor, dt: float) -> typing.Tuple[torch.Tensor, torch.Tensor]:
        """Clamp the control, integrate one step, append to the history (discrete.py:144-167)."""
        assert dt > 0.0
        action = self.make_controls_feasible(action.detach().float().cpu())
        state_new = self.dynamics(self.state_with_time, action, dt).detach().cpu()
        self._state = state_new
        self._history = torch.cat((self._history, state_new.unsqueeze(0)))
        return action, state_new

    def update_inverse(self, state_next: torch.Tensor, dt: float) -> typing.Tuple[torch.Tensor, torch.Tensor]:
        return self.update(self.inverse_dynamics(state_next, self.state, dt), dt)

    def reset(self, state: torch.Tensor, history: torch.Tensor = None):
        assert state.shape == (5,)
        self._state = state.detach().float().cpu()
        self._history = torch.cat((self._history, self._state.unsqueeze(0))) if history is None \
            else torch.as_tensor(history).float().cpu()

    def detach(self):
        self._state, self._history = self._state.detach(), self._history.detach()

    # ---- feasibility -------------------------------------------------------------------------------------------
    def control_limits(self) -> typing.Tuple[float, float]:
        raise NotImplementedError

    def control_norm(self, controls: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def make_controls_feasible(self, controls: torch.Tensor) -> torch.Tensor:
        """Scale the control back into its limits (discrete.py:358-370)."""
        lower, upper = self.control_limits()
        norm = self.control_norm(controls)
        return controls / norm.clamp(min=1e-6) * norm.clamp(lower, upper)

    def check_feasibility_controls(self, controls: torch.Tensor) -> bool:
        lower, upper = self.control_limits()
        norm = self.control_norm(controls)
        return bool(torch.all((lower - 1e-5 <= norm) & (norm <= upper + 1e-5)))

    def check_feasibility_trajectory(self, trajectory: torch.Tensor, dt: float) -> bool:
        return self.check_feasibility_controls(self.roll_trajectory(trajectory, dt))

    # ---- properties --------------------------------------------------------------------------------------------
    @staticmethod
    def expand_state_vector(state_4: torch.Tensor, time: float) -> torch.Tensor:
        return torch.cat((state_4, torch.ones(1) * time))

    @property
    def state(self) -> torch.Tensor:
        return self._state[0:4]

    @property
    def state_with_time(self) -> torch.Tensor:
        return self._state

    @property
    def position(self) -> torch.Tensor:
        return self._state[0:2]

    @property
    def velocity(self) -> torch.Tensor:
        return self._state[2:4]

    @property
    def speed(self) -> float:
        return float(torch.norm(self.velocity))

    @property
    def history(self) -> torch.Tensor:
        return self._history

    @property
    def speed_limits(self) -> typing.Tuple[float, float]:
        v_max = constants.ROBOT_SPEED_MAX if self.is_robot else constants.PED_SPEED_MAX      # discrete.py:543-545
        return -v_max, v_max

    @property
    def is_robot(self) -> bool:
        return self._is_robot

    @property
    def params(self) -> typing.Dict[str, typing.Any]:
        return self._params

    @property
    def color(self) -> np.ndarray:
        return self._color

    @property
    def id(self) -> str:
        return self._id

    @property
    def state_size(self) -> int:
        return 5

    @property
    def control_size(self) -> int:
        return 2

    def __eq__(self, other, check_class: bool = True, check_history: bool = True):
        ok = torch.allclose(self.state_with_time, other.state_with_time) and self.id == other.id
        if check_class:
            ok = ok and self.__class__ == other.__class__
        if check_history:
            ok = ok and self.history.shape == other.history.shape and torch.allclose(self.history, other.history)
        return bool(ok)

    __hash__ = object.__hash__


class LinearDTAgent(DTAgent):
    def __init__(self, position: torch.Tensor, velocity: torch.Tensor = torch.zeros(2), history: torch.Tensor = None,
                 dt: float = None, max_steps: int = None, **agent_kwargs):
        # `dt` / `max_steps` made the reference pre-compute rolling matrices (linear.py:33-48); the CUDA path needs none.
        super().__init__(position, velocity, history=history, **agent_kwargs)


class IntegratorDTAgent(LinearDTAgent):
    ego_kind = "single"

    def control_limits(self) -> typing.Tuple[float, float]:
        v_max = constants.ROBOT_SPEED_MAX if self.is_robot else constants.PED_SPEED_MAX       # integrator_single.py:82-86
        return -v_max, v_max

    def control_norm(self, controls: torch.Tensor) -> torch.Tensor:
        return torch.norm(controls, dim=-1, keepdim=True)                                    # integrator_single.py:88-94

    @property
    def acceleration_max(self) -> float:
        return math.inf                                                                      # integrator_single.py:99-103

    @staticmethod
    def agent_type() -> str:
        return "single_int"


class DoubleIntegratorDTAgent(LinearDTAgent):
    ego_kind = "double"

    def control_limits(self) -> typing.Tuple[float, float]:
        return -self.acceleration_max, self.acceleration_max                                 # integrator_double.py:102-106

    def control_norm(self, controls: torch.Tensor) -> torch.Tensor:
        return torch.abs(controls)                                                           # integrator_double.py:108-114

    @property
    def acceleration_max(self) -> float:
        return constants.ROBOT_ACC_MAX if self.is_robot else constants.PED_ACC_MAX           # discrete.py:547-549

    @staticmethod
    def agent_type() -> str:
        return "double_int"
