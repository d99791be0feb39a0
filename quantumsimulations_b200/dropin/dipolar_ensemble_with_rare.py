"""Drop-in for the reference's ``dipolar_ensemble_with_rare`` module.

Put this directory on ``sys.path`` (``quantumsimulations_b200.dropin_path()``) *before* importing the
reference's ``sweep_sea_detuning.py``: its ``from dipolar_ensemble_with_rare import ...`` (SW:103-109)
then binds to the B200 engine, and the sweep / reprocess / 2D-report scripts run unchanged on top.
Same public names as the reference file; geometry and frequency helpers are numerically identical.
"""
from quantumsimulations_b200.model import (  # noqa: F401
    DipolarRareParams,
    dipolar_couplings_from_positions,
    get_derived_frequencies,
    shell_positions_with_rare_center,
)
from quantumsimulations_b200.api import (  # noqa: F401
    build_hamiltonian_rare,
    initial_state_rare,
    simulate_rare,
    simulate_rare_batch,
)


def dims_with_rare(n_sea: int, is_spin_three_half: bool = False) -> list:
    """Local dimensions (DE:28-34)."""
    return [2] * n_sea + ([4] if is_spin_three_half else [2])
