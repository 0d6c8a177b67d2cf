from .checkpoint_converter import CheckpointConverterBase  # noqa: F401
