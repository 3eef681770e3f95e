"""Free-function twins of the client classes, mirroring the reference's utils/preprocess.py and utils/postprocess.py."""
