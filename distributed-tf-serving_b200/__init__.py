"""B200-native CTR scoring engine: the drop-in for the request path of
neuzxy/Distributed-TF-Serving's Java client (split -> marshal -> score on N replicas ->
ordered merge, DCNClient.java:137-203).

The product is the C-ABI shared library built from ``csrc/`` (``libctr_b200.so``,
declared in ``include/ctr_engine.h``).  This package is the thin host side above it:

* ``binding``     ctypes binding of every entry point of the C ABI (fails loudly when the
                  library is missing: there is no CPU fallback)
* ``dcn_client``  mirror of the reference client's interface (partitionList, sendRequest,
                  processARequest ...) on top of the engine
* ``weights``     writer / reader of the flat weight file the engine loads
* ``build``       nvcc build of the library for sm_100a

The directory name carries a hyphen (it is the name the project was given); import it as
``dtfs_b200`` through the shim module at the repository root.
"""
from . import build, weights  # noqa: F401  (pure python, always importable)
from . import binding, dcn_client  # noqa: F401

__all__ = ["binding", "build", "dcn_client", "weights"]
