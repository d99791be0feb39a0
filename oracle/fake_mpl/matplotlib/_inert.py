"""Inert object: absorbs any attribute access, call, iteration, indexing or `with`.  TEST INFRASTRUCTURE.

matplotlib and tkinter are absent from this image; these stubs only exist so the
reference's sweep/report scripts can be imported and their JSON/NPZ outputs compared.  No figure is
ever produced through them.
"""


class Inert:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return Inert()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return Inert()

    def __getitem__(self, key):
        return Inert()

    def __setitem__(self, key, value):
        pass

    def __iter__(self):
        return iter((Inert(), Inert()))

    def __len__(self):
        return 2

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def __bool__(self):
        return True

    def __float__(self):
        return 0.0

    def flatten(self):
        return [Inert() for _ in range(16)]

    ravel = flatten
