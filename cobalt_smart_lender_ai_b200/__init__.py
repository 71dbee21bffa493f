"""B200-native histogram-GBDT hot path behind the reference's XGBClassifier surface."""
__version__ = "0.1.0"
