"""Writes tests/golden/murmur3_kat.json: PUBLIC known-answer vectors of MurmurHash3_x86_32.

Sources (all public, none from the reference mount, which holds no code or tests):
  * SMHasher's VerificationTest value for MurmurHash3_x86_32, 0xB0F57EE3 (Appleby's
    reference implementation publishes it next to the hash registration);
  * the widely reproduced test vectors for the empty string, short binary keys and the
    ASCII sentences below (seed 0 and seed 0x9747b28c).
The expected values in this file are typed in from those publications; the script only
lays them out as JSON.  tests/test_oracle.py checks the oracle against them, and
tests/test_gpu_hash.py checks the CUDA kernel against them.
"""
import json
import os

VECTORS = [
    # (bytes as hex, seed, expected hash)
    ("", 0x00000000, 0x00000000),
    ("", 0x00000001, 0x514E28B7),
    ("", 0xFFFFFFFF, 0x81F16F39),
    ("ffffffff", 0x00000000, 0x76293B50),
    ("21436587", 0x00000000, 0xF55B516B),
    ("21436587", 0x5082EDEE, 0x2362F9DE),
    ("214365", 0x00000000, 0x7E4A8634),
    ("2143", 0x00000000, 0xA0F7B07A),
    ("21", 0x00000000, 0x72661CF4),
    ("00000000", 0x00000000, 0x2362F9DE),
    ("000000", 0x00000000, 0x85F0B427),
    ("0000", 0x00000000, 0x30F4C306),
    ("00", 0x00000000, 0x514E28B7),
    (b"test".hex(), 0x00000000, 0xBA6BD213),
    (b"Hello, world!".hex(), 0x00000000, 0xC0363E43),
    (b"The quick brown fox jumps over the lazy dog".hex(), 0x00000000, 0x2E4FF723),
    (b"test".hex(), 0x9747B28C, 0x704B81DC),
    (b"Hello, world!".hex(), 0x9747B28C, 0x24884CBA),
    (b"The quick brown fox jumps over the lazy dog".hex(), 0x9747B28C, 0x2FA826CD),
]

out = {
    "algorithm": "MurmurHash3_x86_32",
    "smhasher_verification": 0xB0F57EE3,
    "smhasher_recipe": "for i in 0..255: h[i] = hash(bytes 0..i-1, len i, seed 256-i) little-endian; "
                       "verification = hash(h[0..255] as 1024 bytes, seed 0)",
    "vectors": [{"hex": h, "seed": s, "hash": x} for h, s, x in VECTORS],
}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "murmur3_kat.json")
json.dump(out, open(path, "w"), indent=1)
print(path, len(VECTORS), "vectors")
