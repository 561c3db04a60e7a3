from .hydrometeors import create_hydrometeor
