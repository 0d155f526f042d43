"""tilelang/carver/roller/rasterization.py: the marker type examples compare hints against."""


class Rasterization:
    panel_width_ = None


class NoRasterization(Rasterization):
    pass
