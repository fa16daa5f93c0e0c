"""nls_b200 -- B200-native time-stepping engine for NLSE / sine-Gordon / Klein-Gordon.

Python is only the thin host-side mirror used by tests and bench.py; the product is the
C-ABI library ``libnls_b200.so`` (include/nls.h) and the C++ classes/drivers above it.
"""
from .binding import (  # noqa: F401
    Desc, Engine, NlsError, load_library, library_path, build_library, pinned_empty, pinned_free,
    EQ, INT, LAP, FN, FIELD, RED, REORTH, FLAG_NO_GRAPH, FLAG_NO_PERSISTENT, FLAG_SEWI_HOST2D, FLAG_MARCH, FLAG_TMA, FLAG_PK_SV, FLAG_TMAP, FLAG_NO_TMAP,
    FLAG_DEBUG_PK_FAULT, GEN, GENPOST, GenTerm,
)
from . import slabs  # noqa: F401,E402
from . import fieldgen  # noqa: F401,E402
