"""Per-instance memoisation with a weak reference to ``self``.

Same contract as the reference's helper (src/gemdat/caching.py:7-26): results of
``matrix()``, ``states_prev()``, ``jump_diffusivity()`` ... are computed once per object
and the cache does not keep the object alive.
"""

from __future__ import annotations

import functools
import weakref


def weak_lru_cache(maxsize: int = 128, typed: bool = False):
    def decorate(method):
        @functools.lru_cache(maxsize, typed)
        def cached(ref, *args, **kwargs):
            return method(ref(), *args, **kwargs)

        @functools.wraps(method)
        def bound(self, *args, **kwargs):
            return cached(weakref.ref(self), *args, **kwargs)

        return bound

    return decorate
