class LowDepthAlignmentConfidenceError(Exception):
    """Raised per view when < 1/4 of its SfM points project into the image
    (reference: depth_alignment/exceptions.py:1, points_from_depth.py:124-129)."""
