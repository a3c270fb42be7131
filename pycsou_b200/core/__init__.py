from .linop import LinearOperator
from .functional import Functional, DifferentiableFunctional, ProximableFunctional
from .map import Map, DifferentiableMap, MapStack, DiffMapStack
from .solver import GenericIterativeAlgorithm
