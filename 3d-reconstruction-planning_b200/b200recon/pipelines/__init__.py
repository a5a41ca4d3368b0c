"""`o3d.pipelines` subset: integration (TSDF) and registration (ICP)."""
from . import integration, registration  # noqa: F401
