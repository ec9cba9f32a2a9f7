#!/bin/bash
# Development aid: build a variant of the library next to the product one.
#   tools/build_variant.sh <tag> <extra nvcc flags...>  ->  transmar_gan_b200/libb200mar_<tag>.so
tag=$1; shift
B200MAR_LIB=$(pwd)/transmar_gan_b200/libb200mar_$tag.so B200MAR_NVCC_EXTRA="$*" \
  python -c "from transmar_gan_b200 import _cabi; print(_cabi.build(force=True))"
