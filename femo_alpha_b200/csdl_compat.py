"""`csdl_alpha` if it is installed, else the local stand-in (femo_alpha_b200/csdl_shim.py)."""
try:  # pragma: no cover - csdl_alpha is not installable in the build container
    import csdl_alpha as csdl  # type: ignore
    HAVE_CSDL = True
except Exception:  # ModuleNotFoundError in this image
    from . import csdl_shim as csdl
    HAVE_CSDL = False
