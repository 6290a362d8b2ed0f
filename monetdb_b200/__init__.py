"""B200-native GDK column-at-a-time hot path.

The product is ``libgdk_b200.so`` (hand-written CUDA for sm_100a behind the C-ABI
of ``include/gdk_b200.h``).  This package is the thin Python host mirror of that
ABI used by the tests and ``bench.py``; it never computes anything itself and it
fails loudly when the CUDA library is missing.

Attributes are resolved lazily from :mod:`monetdb_b200.gdk` so that
``python -m monetdb_b200.build`` can run before the library exists.
"""


_SUBMODULES = ("gdk", "pipelines", "shard", "workload", "build")


def __getattr__(name):
    import importlib
    if name in _SUBMODULES:
        # `from monetdb_b200 import workload` must not load libgdk_b200.so (the reference arm of
        # bench.py uses workload / pipelines / shard and must never map the CUDA library)
        return importlib.import_module("." + name, __name__)
    gdk = importlib.import_module(".gdk", __name__)
    try:
        return getattr(gdk, name)
    except AttributeError:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}") from None
