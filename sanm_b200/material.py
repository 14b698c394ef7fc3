"""Material models: graph builders for the stress tensors, mirroring
``fea/material.{h,cpp}`` of the reference (same names, same operator DAGs, so
the lowered device programs are the ones the reference's ``fea`` would run).
"""
from __future__ import annotations

from dataclasses import dataclass

from .capi import SymbolVar

NEOHOOKEAN_I = "neohookean_i"  # incompressible neo-Hookean
NEOHOOKEAN_C = "neohookean_c"  # compressible neo-Hookean
ARAP = "arap"                  # as-rigid-as-possible
STVK_STRETCH = "stvk_stretch"  # stretch term of St. Venant-Kirchhoff


@dataclass
class MaterialProperty:
    """fea/material.h:19-52"""
    young_modulus: float = 0.0
    poisson_ratio: float = 0.0
    bulk_modulus: float = 0.0
    shear_modulus: float = 0.0
    lame_first: float = 0.0
    density: float = 0.0

    @staticmethod
    def from_young_poisson(E: float, nu: float) -> "MaterialProperty":
        # fea/material.cpp:10-18
        return MaterialProperty(
            young_modulus=E, poisson_ratio=nu,
            bulk_modulus=E / (3 * (1 - nu * 2)),
            shear_modulus=E / (2 * (1 + nu)),
            lame_first=E * nu / ((1 + nu) * (1 - nu * 2)))


def pk1(energy_model: str, material: MaterialProperty, F: SymbolVar,
        dim: int = 3) -> SymbolVar:
    """first Piola-Kirchhoff stress, fea/material.cpp:55-99"""
    g = F.graph
    mu = material.shear_modulus
    if energy_model == NEOHOOKEAN_I:
        k = material.bulk_modulus
        FTinv = g.batched_mat_inv_mul(F, None, True).batched_transpose()
        J = F.batched_det()
        Ic = F.pow(2).reduce_sum(-1)
        J23 = J.pow(-2. / 3.)
        t2 = g.linear_combine([(mu / -3., J23 * Ic), (k, J * J), (-k, J)],
                              0) * FTinv
        return g.linear_combine([(mu, J23 * F), (1., t2)])
    if energy_model == NEOHOOKEAN_C:
        lam = material.lame_first
        FTinv = g.batched_mat_inv_mul(F, None, True).batched_transpose()
        J = F.batched_det()
        return g.linear_combine([(mu, F), (-mu, FTinv),
                                 (lam, J.log() * FTinv)])
    if energy_model == ARAP:
        return (F - F.batched_svd_w(True)[2]) * mu
    if energy_model == STVK_STRETCH:
        return g.linear_combine(
            [(mu, F.batched_matmul(F.batched_transpose()).batched_matmul(F)),
             (-mu, F)])
    raise ValueError(f"pk1 unimplemented for energy model {energy_model}")


def cauchy_stress(energy_model: str, material: MaterialProperty, F: SymbolVar,
                  dim: int = 3) -> SymbolVar:
    """Cauchy stress (inverse problems), fea/material.cpp:20-53"""
    g = F.graph
    mu = material.shear_modulus
    if energy_model == NEOHOOKEAN_I:
        k = material.bulk_modulus
        b = F.batched_matmul(F.batched_transpose())
        J = F.batched_det()
        Ic = F.pow(2).reduce_sum(-1)
        J53 = J.pow(-5. / 3.)
        t2 = g.linear_combine([(mu / -3., J53 * Ic), (k, J)],
                              -k).batched_mul_eye(dim)
        return g.linear_combine([(mu, J53 * b), (1., t2)])
    if energy_model == NEOHOOKEAN_C:
        lam = material.lame_first
        b = F.batched_matmul(F.batched_transpose())
        Jinv = F.batched_det().pow(-1)
        xI = g.linear_combine([(mu, Jinv), (lam, Jinv * Jinv.log())])
        return g.linear_combine([(mu, Jinv * b),
                                 (-1., xI.batched_mul_eye(dim))])
    raise ValueError(
        f"cauchy_stress unimplemented for energy model {energy_model}")


def elastic_potential_density(energy_model: str, material: MaterialProperty,
                              F: SymbolVar, dim: int = 3):
    """fea/material.cpp:101-115; None when unsupported"""
    if energy_model == ARAP:
        mu = material.shear_modulus
        return (F - F.batched_svd_w(True)[2]).pow(2).reduce_sum(-1) * (mu / 2)
    return None
