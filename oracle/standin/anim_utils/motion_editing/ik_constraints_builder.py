class IKConstraintsBuilder(object):
    def __init__(self, skeleton, force_orientation=False):
        self.skeleton = skeleton

    def convert_to_ik_constraints(self, constraints, frame_offset=0, time_function=None, constrain_orientation=True):
        return dict()
