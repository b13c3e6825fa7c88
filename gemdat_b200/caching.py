"""Per-instance memoisation of method results.

Same contract as the reference's ``weak_lru_cache`` (src/gemdat/caching.py:7-26) -- results of
``matrix()``, ``states_prev()``, ``jump_diffusivity()`` ... are computed once per object and
the cache never keeps the object alive -- with a different mechanism: the results live in a
small LRU table stored on the instance itself, so they die with it (the reference keys a
module-level ``functools.lru_cache`` on a weak reference to ``self``).
"""

from __future__ import annotations

import functools
from collections import OrderedDict


def weak_lru_cache(maxsize: int = 128, typed: bool = False):
    def decorate(method):
        slot = f'_memo_{method.__name__}'

        @functools.wraps(method)
        def bound(self, *args, **kwargs):
            table: OrderedDict = self.__dict__.setdefault(slot, OrderedDict())
            key = (args, tuple(sorted(kwargs.items())))
            if typed:
                key += tuple(type(a) for a in args)
            if key in table:
                table.move_to_end(key)
                return table[key]
            value = method(self, *args, **kwargs)
            table[key] = value
            if maxsize is not None and len(table) > maxsize:
                table.popitem(last=False)
            return value

        return bound

    return decorate
