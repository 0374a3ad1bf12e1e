"""Runs the UNMODIFIED reference from baseline/_ref (filled in by baseline/install.py)."""


def run(gpus, steps, warmup):
    return {"impl": "reference", "unavailable": "baseline/_ref not built yet"}
