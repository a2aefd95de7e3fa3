"""CPU oracle for the secretstream relay path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``symmetry_b200/`` may import this package; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs do, and only as the checker / the timed CPU arm.

Modules
-------
secretstream  ctypes front-end of ``secretstream_ref.c`` (the C restatement).
libsodium     the real libsodium 1.0.20 bundled with PyNaCl -- the same C
              library the reference reaches through sodium-native 4.1.1
              (/root/reference/package-lock.json:5725-5733).  Used to pin the
              restatement and to generate tests/golden/.
extract       restatement of /root/reference/src/utils.ts:16-52 with
              ECMAScript semantics (no JS engine exists in this image).
framing       restatement of @hyperswarm/secret-stream 6.6.1's uint24-LE
              framing ([RECALLED]).
"""
