"""B200-native batched depth-inference forward path (drop-in for the reference's
``models.model.DepthAnything`` / ``models.enhanced_model.EnhancedDepthAnything``)."""
from .model import DepthAnything, source_scale  # noqa: F401
from .enhanced_model import EnhancedDepthAnything  # noqa: F401
from .refinement import DepthQualityEnhancement, MultiScaleDepthFusion, TextureAwareRefinementModule  # noqa: F401
from .terrain import MarsTerrainMapper, TerrainReconstructor  # noqa: F401
from .layers import disp_to_depth  # noqa: F401
from .pipeline import BatchPipeline  # noqa: F401
from .composition import EnhancedWithRefinement  # noqa: F401
from .runner import run_depth_estimation  # noqa: F401

__version__ = "0.1.0"
