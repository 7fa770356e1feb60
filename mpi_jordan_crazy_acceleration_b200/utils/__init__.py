from .io import MatrixFileError, corner_from_rows, format_corner, read_full, scatter_text_matrix, write_matrix  # noqa: F401
from .report import report_header, report_residual, report_time, usage_line  # noqa: F401
from .timing import ClockSampler  # noqa: F401
