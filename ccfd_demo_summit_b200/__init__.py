"""B200-native random-forest scoring path of the ccfd demo (Seldon model `modelfull`)."""
from .converter import ForestArrays, floor_to_float32, from_arrays, from_sklearn  # noqa: F401

__all__ = ["ForestArrays", "from_sklearn", "from_arrays", "floor_to_float32", "CcfdForestModel", "Forest", "Session"]


def __getattr__(name):
    # model.py needs the CUDA library; import it on first use so the converter works anywhere
    if name in ("CcfdForestModel", "Forest", "Session", "MultiGpuForest", "validate_X"):
        from . import model

        return getattr(model, name)
    raise AttributeError(name)
