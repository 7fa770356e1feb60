from .layout import Layout, find_sender, global_to_local, local_to_global, num_block_rows, owner_of_row, padded_ld, rows_per_process  # noqa: F401
from .comm import Comm  # noqa: F401
