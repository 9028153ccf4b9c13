from .correlation import corr, patchify, CorrLayer, PatchLayer  # noqa: F401
