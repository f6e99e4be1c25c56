"""espnet_b200 -- B200-native CTC hot path (loss fwd+bwd, prefix scoring) behind ESPnet's
own CTC / CTCPrefixScorer interfaces.  See DESIGN.md."""
__version__ = "0.1.0"
