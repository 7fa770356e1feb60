from .eager import EagerJordan, formula_rows, minloc_fold  # noqa: F401
from .jordan import FORMULAS, JordanSolver, invert  # noqa: F401
