"""Planner configuration (the one dataclass a maintainer adds to ``TyGrit/planning/config.py``).

Field names shared with ``VampPlannerConfig`` (``TyGrit/planning/config.py:9-47``) keep their
meaning and defaults so an existing TOML section keeps working; the B200-specific fields are new.
"""

from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class PlannerConfig:
    timeout: float = 5.0
    point_radius: float = 0.03


@dataclass(frozen=True)
class B200PlannerConfig(PlannerConfig):
    # shared with VampPlannerConfig
    algorithm: str = "rrtc"
    whole_body_algorithm: str = "multilayer_rrtc"
    sampler: str = "halton"
    simplify: bool = True
    interpolation_steps: int = 16
    interpolation_density: float = 0.08
    bounds_padding: float = 0.1
    fcit_max_iterations: int = 200
    fcit_max_samples: int = 8192
    fcit_batch_size: int = 4096
    fcit_reverse_weight: float = 10.0
    fcit_optimize: bool = True
    # B200 engine
    device: int = 0
    max_points: int = 1 << 20
    voxels_per_metre: float = 50.0
    rake: int = 8                 # VAMP's SIMD rake for Fetch (oracle assumption 6)
    resolution: float = 32.0      # samples per unit of configuration-space distance
    rrt_batch: int = 512          # candidate extensions validated per iteration
    rrt_range: float = 1.0        # maximum extension length (configuration-space L2)
    rrt_max_iterations: int = 400
    shortcut_rounds: int = 4
    seed: int = 0
