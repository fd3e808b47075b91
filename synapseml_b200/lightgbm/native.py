"""ctypes binding of lib_lightgbm.so — the stand-in for the SWIG class `com.microsoft.ml.lightgbm.lightgbmlib`.

The reference loads lib_lightgbm.so through NativeLoader (LightGBMUtils.initializeNativeLibrary,
lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/LightGBMUtils.scala:31-35) and turns a -1 return code
into an exception carrying LGBM_GetLastError() (LightGBMUtils.validate, :15-20).  Same here.  There is no fallback: if
the CUDA library is missing or cannot be loaded, importing this module's `lightgbmlib()` raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(os.path.dirname(_HERE), "lib", "lib_lightgbm.so")
_lib = None

C_API_DTYPE_FLOAT32, C_API_DTYPE_FLOAT64, C_API_DTYPE_INT32, C_API_DTYPE_INT64 = 0, 1, 2, 3


class LightGBMNativeError(RuntimeError):
    pass


def library_path():
    return _SO


def lightgbmlib():
    """Loads (once) and returns the native library handle."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise LightGBMNativeError(
                "%s is missing: build it with `python -m synapseml_b200.build` (there is no CPU fallback)" % _SO)
        lib = C.CDLL(_SO)
        lib.LGBM_GetLastError.restype = C.c_char_p
        _lib = lib
    return _lib


def validate(rc, component="LightGBM"):
    """LightGBMUtils.validate (LightGBMUtils.scala:15-20)."""
    if rc == -1:
        msg = lightgbmlib().LGBM_GetLastError().decode(errors="replace")
        raise LightGBMNativeError("%s call failed in LightGBM with error: %s" % (component, msg))
    return rc
