"""o3d.pipelines look-alike: integration (TSDF), odometry, registration."""
from . import integration, odometry, registration, sharded  # noqa: F401
