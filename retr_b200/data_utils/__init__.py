"""GPU counterparts of the reference's ``data_utils`` pieces that sit immediately before the hot path."""
from .preprocess import preprocess_batch  # noqa: F401
