from ._inert import Inert


def subplots(nrows=1, ncols=1, *a, **k):
    return Inert(), Inert()


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return Inert()
