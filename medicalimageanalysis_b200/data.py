"""Minimal stand-in for the reference's global registry (``data.py:59-95``): class-level dicts that let
``Rigid`` look images up by name, exactly as ``structure/rigid.py:666-672,689`` does.  Only the image and
rigid tables are on the hot path; ROI / POI / dose / deformable tables are out of scope."""


class Data(object):
    image = {}
    rigid = {}
    image_list = []
    rigid_list = []
    deformable_list = []          # names of displacement fields read so far (read/mhd.py:222-233 resolves collisions on it)

    @classmethod
    def clear(cls):
        """``data.py:71-95``."""
        cls.image = {}
        cls.rigid = {}
        cls.image_list = []
        cls.rigid_list = []
        cls.deformable_list = []

    @classmethod
    def delete_image(cls, image_name):
        if image_name in cls.image:
            del cls.image[image_name]
            cls.image_list.remove(image_name)
