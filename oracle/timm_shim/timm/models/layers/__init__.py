from .helpers import to_2tuple


def __getattr__(name):  # lazy: avoids the vision_transformer <-> layers import cycle
    if name == "PatchEmbed":
        from ..vision_transformer import PatchEmbed
        return PatchEmbed
    raise AttributeError(name)
