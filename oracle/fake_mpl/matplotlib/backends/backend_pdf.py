from .._inert import Inert


class PdfPages(Inert):
    """Accepts a path, writes nothing."""
