"""DroneEnvParams / Action / Object -- the reference's config and enums.

Mirrors /root/reference/jax_impl/env/env.py:11-26 (field names, order and defaults) and
/root/reference/common/constants.py:3-19.
"""
from dataclasses import dataclass, replace as _replace
from enum import IntEnum


class Action(IntEnum):
    LEFT = 0
    DOWN = 1
    RIGHT = 2
    UP = 3
    STAY = 4

    @classmethod
    def num_actions(cls) -> int:
        return len(cls)


class Object(IntEnum):
    SKYSCRAPER = 2
    STATION = 3
    DROPZONE = 4
    PACKET = 5


@dataclass(frozen=True)
class DroneEnvParams:
    grid_size: int = 8
    n_drones: int = 3
    pickup_reward: float = 0.0
    delivery_reward: float = 1.0
    crash_reward: float = -1.0
    charge_reward: float = -0.1
    discharge: int = 10
    charge: int = 20
    packets_factor: int = 3
    dropzones_factor: int = 2
    stations_factor: int = 2
    skyscrapers_factor: int = 3
    wrapper: str = 'window'
    window_radius: int = 3

    def replace(self, **updates):
        """flax.struct.dataclass-style functional update."""
        return _replace(self, **updates)

    def validate(self):
        """The ValueErrors of reset, jax_impl/env/env.py:91-104."""
        num_packets = self.packets_factor * self.n_drones
        num_dropzones = self.dropzones_factor * self.n_drones
        num_stations = self.stations_factor * self.n_drones
        num_skyscrapers = self.skyscrapers_factor * self.n_drones
        num_objects = num_packets + num_skyscrapers + num_dropzones + num_stations
        if num_objects > self.grid_size ** 2:
            raise ValueError(
                f'Grid supports only {self.grid_size**2:,} positions but {num_objects:,} objects '
                f'({num_dropzones:,} dropzones, {num_stations:,} charging stations, {num_packets:,} packages, '
                f'{num_skyscrapers:,} skyscrapers) were attempted to be placed.')
        elif self.n_drones > self.grid_size ** 2:
            raise ValueError(
                f'Grid supports only {self.grid_size**2:,} positions but {self.n_drones:,} drones were '
                'attempted to be placed.')

    @property
    def window_size(self) -> int:
        return 2 * self.window_radius + 1

    @property
    def obs_size(self) -> int:
        """Flattened per-drone observation length, W*W*6 (train_jax.py:174)."""
        return self.window_size ** 2 * 6

    def algorithmic_bytes_per_env_step(self) -> int:
        """24 n W^2 + 2 G^2 + 14 n  (SURVEY.md 8d)."""
        return 24 * self.n_drones * self.window_size ** 2 + 2 * self.grid_size ** 2 + 14 * self.n_drones
