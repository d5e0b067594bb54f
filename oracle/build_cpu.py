"""TEST INFRASTRUCTURE — compiles the plain-C restatements under oracle/ (currently cpu_raster.c)."""
from __future__ import annotations


def build() -> str:
    from oracle import cpu_raster
    return cpu_raster.build()


if __name__ == "__main__":
    print(build())
