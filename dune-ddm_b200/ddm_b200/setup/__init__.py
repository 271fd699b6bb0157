"""Host-side setup restatement (runs once; not on the solve-phase hot path)."""
from . import coarse, decomp, factor, fem, mesh, problem  # noqa: F401
