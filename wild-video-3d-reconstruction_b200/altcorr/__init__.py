from .correlation import corr, corr_stacked, patchify, CorrLayer, PatchLayer  # noqa: F401
