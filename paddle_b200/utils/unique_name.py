"""Unique name generator (reference: python/paddle/utils/unique_name.py, base/unique_name.py).
Parameter names (``linear_0.w_0``) leak into optimizer state keys and sharding order, so the scheme is the same:
a per-prefix counter, ``generate(key) -> f"{key}_{n}"``, plus ``guard`` to swap generators."""
from __future__ import annotations

import contextlib
from collections import defaultdict


class UniqueNameGenerator:
    def __init__(self, prefix: str = ""):
        self.ids = defaultdict(int)
        self.prefix = prefix or ""

    def __call__(self, key: str) -> str:
        n = self.ids[key]
        self.ids[key] += 1
        return f"{self.prefix}{key}_{n}"


generator = UniqueNameGenerator()


def generate(key: str) -> str:
    return generator(key)


def switch(new_generator=None):
    global generator
    old = generator
    generator = new_generator if new_generator is not None else UniqueNameGenerator()
    return old


@contextlib.contextmanager
def guard(new_generator=None):
    if isinstance(new_generator, str):
        new_generator = UniqueNameGenerator(new_generator)
    old = switch(new_generator)
    try:
        yield
    finally:
        switch(old)
