"""Mirror of the hot-path parts of the reference's tools/ package (YOLOv4 head decode + numpy NMS)."""
