from . import p2p_communication  # noqa: F401
