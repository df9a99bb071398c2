from .pp_layers import LayerDesc, PipelineLayer, SegmentLayers, SharedLayerDesc  # noqa: F401
