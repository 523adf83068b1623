"""lua-mapreduce-1_b200 -- B200-native data plane behind lua-mapreduce's task-module API.

Import as `lua_mapreduce_1_b200` (the directory name has a hyphen; the sibling
`lua_mapreduce_1_b200/` package points its __path__ here).

  _native     ctypes binding of the C ABI (include/mrgpu.h -> libmrgpu.so)
  mapreduce   host-side mirror of the reference's server / worker / task-module API
  corpus      synthetic Europarl-shaped corpus generator
  build       in-tree nvcc / gcc build
"""
__version__ = "0.1.0"
