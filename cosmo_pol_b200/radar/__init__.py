from .pyart_wrapper import RadopVolume, PyartRadop  # noqa: F401
