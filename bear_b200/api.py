"""`visr_bear.api`-shaped Python surface of the CUDA render core.

The classes come from the pybind11 module `bear_b200._api` (bear_b200/csrc/pyapi.cpp), which wraps
include/bear_b200/host.hpp the way visr_bear/pythonwrappers/api.cpp:23-181 wraps bear/api.hpp: Config,
Time, Listener, ObjectsInput, DirectSpeakersInput, HOAInput (+ HOATypeMetadata), Renderer with
add_*_block / process / get_block_start_time / set_block_start_time / set_listener, and
DataFileMetadata.read_from_file.  ObjectsInput / DirectSpeakersInput carry the gain vectors libear's
gain calculators would produce (`direct_gains`, `diffuse_gains`, `gains`) instead of ADM type metadata:
libear needs Boost and is outside this repository."""
try:
    from ._api import (Config, DataFileMetadata, DirectSpeakersInput, HOAInput, HOATypeMetadata, Listener,  # noqa: F401
                       MetadataInput, ObjectsInput, Renderer, Time)
except ImportError as e:   # not built: there is no fallback
    raise ImportError("bear_b200._api is not built: run `python -c 'import __graft_entry__ as g; g.build()'` (%s)" % e)
