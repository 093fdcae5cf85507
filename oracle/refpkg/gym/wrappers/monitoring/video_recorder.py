"""Import stub: the reference's RecordingSafeLifeWrapper subclasses this; never instantiated here."""


class VideoRecorder(object):
    def __init__(self, *args, **kwargs):
        raise NotImplementedError("video recording is out of scope (no graphics stack)")


class ImageEncoder(object):
    def __init__(self, *args, **kwargs):
        raise NotImplementedError("video recording is out of scope (no graphics stack)")
