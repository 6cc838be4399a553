"""B200-native ProtoSeg prototype head + push search (drop-in for the reference's model.PPNet /
segmentation.push.push_prototypes hot path).  See DESIGN.md."""
__version__ = "0.1.0"
