class BVHReader(object):
    def __init__(self, filename):
        raise NotImplementedError("BVH I/O is outside the hot path (SURVEY.md section 2 #16)")
