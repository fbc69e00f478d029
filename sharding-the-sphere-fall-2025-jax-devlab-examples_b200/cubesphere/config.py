"""Configuration: the reference's ``parallelization`` block plus the solver keys this build adds.

The reference reads ``self.config['parallelization']`` with three keys and these defaults
(JAX-DevLab-Examples.py:21-24, deck slide 8):

    parallelization:
      tiles_per_edge: 1      # 1 -> 6 tiles, 2 -> 24, 3 -> 54
      num_devices: 6         # <= 6 * tiles_per_edge^2 and a divisor of it
      device_type: 'cpu'     # 'gpu' or 'cpu' (cpu = host-side planning / tests only here)

The key names and defaults are kept verbatim.  ``solver`` is new (the reference ships no solver).
"""
from __future__ import annotations

from copy import deepcopy

DEFAULTS = {
    "parallelization": {"tiles_per_edge": 1, "num_devices": 6, "device_type": "cpu"},
    "solver": {"N": 48, "halo": 2, "dtype": "f64", "rk": "ssprk3", "limiter": "mc", "case": "tc2", "cfl": 0.5},
}


def load_config(source=None):
    """``source``: None (defaults), a dict, or a path to a YAML file.  Unknown sections are kept,
    unknown keys inside the two known sections raise ValueError (typos should not pass silently)."""
    cfg = deepcopy(DEFAULTS)
    if source is None:
        return cfg
    if isinstance(source, (str, bytes)) or hasattr(source, "__fspath__"):
        import yaml
        with open(source) as fh:
            source = yaml.safe_load(fh) or {}
    if not isinstance(source, dict):
        raise ValueError("configuration must be a mapping")
    for section, values in source.items():
        if section in DEFAULTS:
            if not isinstance(values, dict):
                raise ValueError(f"section '{section}' must be a mapping")
            unknown = set(values) - set(DEFAULTS[section])
            if unknown:
                raise ValueError(f"unknown key(s) in '{section}': {sorted(unknown)}")
            cfg[section].update(values)
        else:
            cfg[section] = deepcopy(values)
    return cfg
