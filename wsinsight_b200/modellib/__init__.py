"""Model loading, fused data path and inference loop (mirror of ``wsinsight/modellib``)."""
