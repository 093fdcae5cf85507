"""Stand-in for gym's VideoRecorder, which the reference's SafeLifeRecorder subclasses
(env_wrappers.py:95-136).  It renders nothing (no renderer on this box): only the attributes and the
no-op methods SafeLifeRecorder relies on, so that the reference's trajectory .npz can be generated for
the golden fixtures."""


class VideoRecorder(object):
    def __init__(self, env, path=None, metadata=None, enabled=True, base_path=None):
        self.env = env
        self.enabled = enabled
        self.path = path if path is not None else (base_path or "video") + ".mp4"

    def capture_frame(self):
        pass

    def close(self):
        pass


class ImageEncoder(object):
    def __init__(self, *args, **kwargs):
        raise NotImplementedError("video recording is out of scope (no graphics stack)")
