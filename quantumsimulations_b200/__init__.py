"""B200-native time-evolution engine for the dipolar sea + rare spin-1/2 ensemble.

Drop-in for the one hot path of TimHarrelson/QuantumSimulations: ``simulate_rare`` ->
``build_hamiltonian_rare`` -> ``qt.sesolve`` (dipolar_ensemble_with_rare.py:611, :453, :653).
See DESIGN.md / INTEGRATION.md.  The CUDA library (libqsb200.so) is mandatory: no CPU fallback.
"""
import os as _os

from .model import (  # noqa: F401
    DipolarRareParams,
    dipolar_couplings_from_positions,
    get_derived_frequencies,
    shell_positions_with_rare_center,
)
from .table import SpinTable, compile_table  # noqa: F401
from .api import (  # noqa: F401
    SpinOperator,
    build_hamiltonian_rare,
    initial_state_rare,
    simulate_rare,
    simulate_rare_batch,
)

__version__ = "0.1.0"


def dropin_path() -> str:
    """Directory to prepend to sys.path so that ``import dipolar_ensemble_with_rare`` resolves to the engine."""
    return _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "dropin")
