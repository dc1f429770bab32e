from .adaptive_pipeline import AdaptivePipeline, AdaptiveRouter, ContentAnalysis, RoutingStats  # noqa: F401
