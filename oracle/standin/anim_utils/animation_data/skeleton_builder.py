from .skeleton import Skeleton


class SkeletonBuilder(object):
    def load_from_bvh(self, bvh_reader, animated_joints=None):
        raise NotImplementedError("BVH I/O is outside the hot path")

    def load_from_desc(self, desc):
        return Skeleton.from_desc(desc)
