"""Platform plugin modules.

The reference ships ``pyopal.platform.{sse2,sse4,avx2,neon}``
(/root/reference/src/pyopal/platform/CMakeLists.txt:3-32); this package ships
exactly one: ``cuda`` (sm_100a).  No multi-backend dispatch, no CPU fallback.
"""
