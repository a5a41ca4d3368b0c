"""`o3d.pipelines` subset: integration (TSDF), registration (ICP) and the fused ICP -> integrate -> mesh path."""
from . import integration, registration, fusion  # noqa: F401
