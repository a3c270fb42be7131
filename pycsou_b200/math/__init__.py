from .prox import *
