from .base import *
from .conv import *
from .diff import *
from .sampling import *
from .transform import *
