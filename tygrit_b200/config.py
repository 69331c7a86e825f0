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
    self_collision: bool = True   # VAMP's verdict includes self-collision; pairs = rl/train.py:78-107 SELF_COLLISION_PAIRS
    rrt_batch: int = 512          # candidate extensions validated per iteration
    rrt_range: float = 1.0        # maximum extension length (configuration-space L2)
    rrt_max_iterations: int = 400
    rrt_max_nodes: int = 65536    # capacity of each device-resident tree
    shortcut_rounds: int = 2      # rounds of all-pairs shortcutting (each: one batched edge check + a DAG shortest path)
    seed: int = 0                 # offset of the Halton sequence
    # base layer of multilayer_rrtc (algorithms.py:93-96: HybridAStarConfig.reverse_penalty = 50)
    reverse_penalty: float = 50.0
    base_resolution: float = 0.1  # lattice spacing [m] (16 headings)
    base_turn_cost: float = 0.3   # metres of path cost per radian turned in place
    base_max_rounds: int = 16     # lazy edge-validation rounds
    fcit_max_edge: float = 4.0    # the roadmap is complete up to this configuration-space edge length
