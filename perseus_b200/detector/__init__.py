from .models import KeypointCNN  # noqa: F401
from .decode import denormalize_pixel_coordinates, normalize_pixel_coordinates  # noqa: F401
from .streaming import StreamingDetector  # noqa: F401
from .pipeline import BatchPipeline  # noqa: F401
