"""Native (C++) host runtime of the inference path: safetensors parallel loader and bucket router."""
from .native import NativeRouter, SafetensorsFile, available, build, load_safetensors  # noqa: F401
