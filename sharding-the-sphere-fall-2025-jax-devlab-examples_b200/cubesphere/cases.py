"""Initial conditions: Williamson et al. (1992) cases 1, 2, 5, 6 and Galewsky et al. (2004).

None of this is in the reference (its "Initial Conditions: Physics" box, deck
slide 6, is not shipped; the cosine bell of slides 13/18 -- peak ~1000 at lon
270, lat 0 on face 3 -- is Williamson case 1).  Formulas are from the papers.

Every case returns a :class:`Case` holding interior fields ``h, u1, u2`` of shape
``[6, N, N]`` (fp64; u1, u2 are panel-local *contravariant* velocity components
in rad/s), the topography ``hs`` (or None), the rotation axis used for the
Coriolis parameter and whether the velocity is frozen (pure advection).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import geometry as geo

DAY = 86400.0


@dataclass
class Case:
    name: str
    N: int
    h: np.ndarray
    u1: np.ndarray
    u2: np.ndarray
    hs: np.ndarray | None = None
    axis: tuple = (0.0, 0.0, 1.0)
    advect_only: bool = False
    meta: dict = field(default_factory=dict)


def _solid_body_wind(lon, lat, u0, alpha0):
    u = u0 * (np.cos(lat) * np.cos(alpha0) + np.sin(lat) * np.cos(lon) * np.sin(alpha0))
    v = -u0 * np.sin(lon) * np.sin(alpha0)
    return u, v


def _to_panel(u_east, v_north, pts, patch, a):
    V = geo.cartesian_from_lonlat_wind(u_east, v_north, pts)
    return geo.contravariant_from_cartesian(V, patch.N, a, patch=patch)


def tc1(N, alpha0=0.0, a=geo.EARTH_RADIUS, p=1, tiles=None):
    """Cosine-bell advection by solid-body rotation (velocity frozen)."""
    patch = geo.Patch(N, p, tiles)
    pts = patch.points()
    lon, lat = geo.lonlat(pts)
    u0 = 2.0 * np.pi * a / (12.0 * DAY)
    ue, vn = _solid_body_wind(lon, lat, u0, alpha0)
    u1, u2 = _to_panel(ue, vn, pts, patch, a)
    lon_c, lat_c, h0, R = 1.5 * np.pi, 0.0, 1000.0, a / 3.0
    cosd = np.sin(lat_c) * np.sin(lat) + np.cos(lat_c) * np.cos(lat) * np.cos(lon - lon_c)
    r = a * np.arccos(np.clip(cosd, -1.0, 1.0))
    h = np.where(r < R, 0.5 * h0 * (1.0 + np.cos(np.pi * r / R)), 0.0)
    return Case("tc1", N, h, u1, u2, None, (-np.sin(alpha0), 0.0, np.cos(alpha0)), True,
                dict(u0=u0, alpha0=alpha0))


def _geostrophic_height(lon, lat, gh0, u0, alpha0, a, omega, grav):
    s = -np.cos(lon) * np.cos(lat) * np.sin(alpha0) + np.sin(lat) * np.cos(alpha0)
    return (gh0 - (a * omega * u0 + 0.5 * u0 * u0) * s * s) / grav


def tc2(N, alpha0=0.0, a=geo.EARTH_RADIUS, omega=geo.EARTH_OMEGA, grav=geo.EARTH_GRAVITY, p=1, tiles=None):
    """Steady zonal geostrophic flow; the initial condition is the exact solution."""
    patch = geo.Patch(N, p, tiles)
    pts = patch.points()
    lon, lat = geo.lonlat(pts)
    u0 = 2.0 * np.pi * a / (12.0 * DAY)
    ue, vn = _solid_body_wind(lon, lat, u0, alpha0)
    u1, u2 = _to_panel(ue, vn, pts, patch, a)
    h = _geostrophic_height(lon, lat, 2.94e4, u0, alpha0, a, omega, grav)
    return Case("tc2", N, h, u1, u2, None, (-np.sin(alpha0), 0.0, np.cos(alpha0)), False,
                dict(u0=u0, alpha0=alpha0))


def tc5(N, a=geo.EARTH_RADIUS, omega=geo.EARTH_OMEGA, grav=geo.EARTH_GRAVITY, p=1, tiles=None):
    """Zonal flow over an isolated conical mountain."""
    patch = geo.Patch(N, p, tiles)
    pts = patch.points()
    lon, lat = geo.lonlat(pts)
    u0, h0 = 20.0, 5960.0
    ue, vn = _solid_body_wind(lon, lat, u0, 0.0)
    u1, u2 = _to_panel(ue, vn, pts, patch, a)
    total = _geostrophic_height(lon, lat, grav * h0, u0, 0.0, a, omega, grav)
    R, lon_c, lat_c = np.pi / 9.0, 1.5 * np.pi, np.pi / 6.0
    r = np.sqrt(np.minimum(R * R, (lon - lon_c) ** 2 + (lat - lat_c) ** 2))
    hs = 2000.0 * (1.0 - r / R)
    return Case("tc5", N, total - hs, u1, u2, hs, (0.0, 0.0, 1.0), False, dict(u0=u0))


def tc6(N, a=geo.EARTH_RADIUS, omega=geo.EARTH_OMEGA, grav=geo.EARTH_GRAVITY, p=1, tiles=None):
    """Rossby-Haurwitz wave, wavenumber 4."""
    patch = geo.Patch(N, p, tiles)
    pts = patch.points()
    lon, lat = geo.lonlat(pts)
    w = K = 7.848e-6
    R, h0 = 4.0, 8000.0
    c = np.cos(lat)
    s = np.sin(lat)
    ue = a * w * c + a * K * c ** (R - 1) * (R * s * s - c * c) * np.cos(R * lon)
    vn = -a * K * R * c ** (R - 1) * s * np.sin(R * lon)
    A = 0.5 * w * (2 * omega + w) * c * c + 0.25 * K * K * c ** (2 * R) * (
        (R + 1) * c * c + (2 * R * R - R - 2) - 2 * R * R / np.maximum(c * c, 1e-300))
    B = 2 * (omega + w) * K / ((R + 1) * (R + 2)) * c ** R * ((R * R + 2 * R + 2) - (R + 1) ** 2 * c * c)
    C = 0.25 * K * K * c ** (2 * R) * ((R + 1) * c * c - (R + 2))
    h = h0 + (a * a / grav) * (A + B * np.cos(R * lon) + C * np.cos(2 * R * lon))
    u1, u2 = _to_panel(ue, vn, pts, patch, a)
    return Case("tc6", N, h, u1, u2, None, (0.0, 0.0, 1.0), False, {})


def galewsky(N, perturbed=True, a=geo.EARTH_RADIUS, omega=geo.EARTH_OMEGA, grav=geo.EARTH_GRAVITY, p=1, tiles=None):
    """Barotropically unstable mid-latitude jet (Galewsky, Scott & Polvani 2004)."""
    patch = geo.Patch(N, p, tiles)
    pts = patch.points()
    lon, lat = geo.lonlat(pts)
    umax, phi0 = 80.0, np.pi / 7.0
    phi1 = 0.5 * np.pi - phi0
    en = np.exp(-4.0 / (phi1 - phi0) ** 2)

    def jet(phi):
        inside = (phi > phi0) & (phi < phi1)
        d = np.where(inside, (phi - phi0) * (phi - phi1), -1.0)
        return np.where(inside, umax / en * np.exp(1.0 / d), 0.0)

    # balanced height by quadrature on a fine latitude grid (composite Simpson via midpoints)
    nq = 20001
    phi = np.linspace(-0.5 * np.pi, 0.5 * np.pi, nq)
    uq = jet(phi)
    integrand = a * uq * (2.0 * omega * np.sin(phi) + np.tan(np.clip(phi, -1.5, 1.5)) * uq / a)
    integ = np.concatenate([[0.0], np.cumsum(0.5 * (integrand[1:] + integrand[:-1]) * np.diff(phi))])
    hq = -integ / grav
    mean = np.sum(0.5 * (hq[1:] * np.cos(phi[1:]) + hq[:-1] * np.cos(phi[:-1])) * np.diff(phi)) / 2.0
    hq = hq + (10000.0 - mean)
    h = np.interp(lat, phi, hq)
    if perturbed:
        lam = np.where(lon > np.pi, lon - 2.0 * np.pi, lon)
        h = h + 120.0 * np.cos(lat) * np.exp(-(lam * 3.0) ** 2) * np.exp(-((np.pi / 4.0 - lat) * 15.0) ** 2)
    u1, u2 = _to_panel(jet(lat), np.zeros_like(lat), pts, patch, a)
    return Case("galewsky", N, h, u1, u2, None, (0.0, 0.0, 1.0), False, {})


CASES = {"tc1": tc1, "tc2": tc2, "tc5": tc5, "tc6": tc6, "galewsky": galewsky}


def make_case(name: str, N: int, **kw) -> Case:
    try:
        fn = CASES[name]
    except KeyError:
        raise ValueError(f"Unknown case: {name}; choose from {sorted(CASES)}") from None
    case = fn(N, **kw)
    case.meta["patch"] = geo.Patch(N, kw.get("p", 1), kw.get("tiles"))
    return case


def stable_dt(case: Case, cfl=0.5, a=geo.EARTH_RADIUS, grav=geo.EARTH_GRAVITY):
    """Time step from the advective + gravity-wave CFL number in (alpha, beta) space."""
    N = case.N
    patch = case.meta.get("patch") or geo.Patch(N)
    m = geo.metric_terms(patch.X, patch.Y, a)
    c = 0.0 if case.advect_only else np.sqrt(grav * np.maximum(case.h + (0 if case.hs is None else 0), 0.0))
    s1 = np.abs(case.u1) + c * np.sqrt(m["gi11"])
    s2 = np.abs(case.u2) + c * np.sqrt(m["gi22"])
    delta = 0.5 * np.pi / N
    return float(cfl * delta / max(s1.max(), s2.max()))
