from .align import align_video, test_alignment, window_plan, roc_auc  # noqa: F401
