"""Fused-iteration planner (filled in as fused kernels land)."""


def plan_pds(solver):
    return None


def plan_apgd(solver):
    return None
