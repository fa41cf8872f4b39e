"""B200-native (sm_100a) DDIM sampling path of DDIM-COLD -- a drop-in for the reference's
``ViT.py`` / ``ViT_draft2drawing.py`` model class.  See DESIGN.md and INTEGRATION.md."""
from .vit import DiffusionVisionTransformer
from .checkpoint import infer_shape, load_checkpoint, model_from_yaml
from .shard import SharedHostImages, shard_bounds, sample_sharded
from .evaluate import evaluate, smooth_l1_loss

__all__ = ["DiffusionVisionTransformer", "infer_shape", "load_checkpoint", "model_from_yaml",
           "shard_bounds", "sample_sharded", "SharedHostImages", "evaluate", "smooth_l1_loss"]
