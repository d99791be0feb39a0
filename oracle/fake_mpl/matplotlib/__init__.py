from ._inert import Inert

__version__ = "0.0-stub"
rcParams = {}


def use(*a, **k):
    return None


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return Inert()
