from .base import *
from .penalty import *
from .loss import *
