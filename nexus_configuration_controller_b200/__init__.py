"""B200-native drift check for nexus-configuration-controller (one hot path).

``arena`` lays informer-cache contents out for the GPU, ``drift`` drives libncdrift.so through
its C ABI (include/ncdrift.h).  Importing the package does not load the library; the first
``drift.Context`` does, and fails loudly when it is missing or no B200 is visible.
"""
from . import arena  # noqa: F401

__all__ = ["arena", "drift"]
