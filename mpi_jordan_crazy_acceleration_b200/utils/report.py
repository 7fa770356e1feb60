"""The reference's stdout report, byte-compatible:

    A\\n <corner> glob_time: %.2f\\n inverse matrix:\\n\\n <corner> residual: %e\\n

Error lines: ``usage:%s n m [<file>]`` (rc 1), ``cannot open %s`` / ``cannot read %s`` /
``singular matrix`` / ``Not enough memory!`` (rc 2).  Rank 0 prints, everybody returns the same rc.
"""
from __future__ import annotations


def c_format(fmt: str, value: float) -> str:
    """printf-exact formatting of one double (Python's % operator matches C for %e / %.2f)."""
    return fmt % value


def usage_line(argv0: str) -> str:
    return "usage:%s n m [<file>]\n" % argv0


def report_header() -> str:
    return "A\n"


def report_time(seconds: float) -> str:
    return "glob_time: %.2f\n" % seconds + "inverse matrix:\n\n"


def report_residual(res: float) -> str:
    return "residual: %e\n" % res

