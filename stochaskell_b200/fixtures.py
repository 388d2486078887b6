"""Model inputs shared by tests and bench (no reference access at run time)."""
import os

import numpy as np

_GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

COAL_T = 40907.0  # app/rj/Main.hs:228


def coal_days() -> np.ndarray:
    """coalData' (app/rj/Main.hs:109-111): days since the first disaster, 191 ascending doubles."""
    yr = np.loadtxt(os.path.join(_GOLD, "coal_years.txt"))
    days = 365.2425 * yr
    return days - days[0]
