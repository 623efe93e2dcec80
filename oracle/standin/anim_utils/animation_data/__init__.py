from .skeleton import Skeleton, SkeletonNode
