"""ref-geotransformer_b200: B200-native (sm_100a) GeoTransformer geometric self/cross-attention hot path.

Drop-in classes for /root/reference/geotransformer/modules/geotransformer (GeometricTransformer,
GeometricStructureEmbedding) and /root/reference/geotransformer/modules/transformer (RPEConditionalTransformer,
RPETransformerLayer, TransformerLayer, SinusoidalPositionalEmbedding, ...).  Import via pkgpath.load_package()."""
from . import _abi, dist, ops, synth  # noqa: F401
from .modules import (  # noqa: F401
    AttentionLayer, AttentionOutput, GeometricStructureEmbedding, GeometricTransformer, GraphedGeometricTransformer, PipelinedGeometricTransformer,
    MultiHeadAttention,
    RPEAttentionLayer, RPEConditionalTransformer, RPEMultiHeadAttention, RPETransformerLayer,
    SinusoidalPositionalEmbedding, TransformerLayer, build_act_layer, build_dropout_layer)

__all__ = [
    'GeometricTransformer', 'GraphedGeometricTransformer', 'PipelinedGeometricTransformer', 'GeometricStructureEmbedding', 'RPEConditionalTransformer', 'RPETransformerLayer',
    'RPEAttentionLayer', 'RPEMultiHeadAttention', 'TransformerLayer', 'AttentionLayer', 'MultiHeadAttention',
    'AttentionOutput', 'SinusoidalPositionalEmbedding', 'ops', 'synth',
]
