#!/usr/bin/env bash
# Builds the UNMODIFIED reference (ddkang/avrecode at /root/reference) into oracle/_ref/.
# Test infrastructure only: the binaries are the parity oracle and the timed CPU baseline.
# Nothing from /root/reference is copied into the repository; all outputs land in
# oracle/_ref/ (git-ignored).  Recipe = SURVEY.md section 8(c):
#   1. out-of-tree configure+make of the vendored libavcodec-hooks fork (h264 decoder only,
#      portable C bin decoder: --disable-inline-asm keeps every bin on the hook path);
#   2. protobuf replaced by oracle/ref_shim/recode.pb.h (protoc is not installed);
#   3. recode.cpp compiled from a sed-patched temporary copy (GCC 13 rejects the
#      std::function<void(T*&)> unique_ptr deleter, recode.cpp:34-48) -- semantics unchanged.
# Variants:
#   recode_ref      -O2 -DNDEBUG         timing baseline
#   recode_ref_dbg  -O2, asserts on      correctness oracle
#   recode_ref_h8   -O2 -DNDEBUG + the one-line decompressor fix for 8x8 blocks
#   recode_ref_trace like _dbg, prints every coded bin (key, symbol) to fd 3 when AVR_TRACE=1
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
B="$OUT/build"
JOBS=${JOBS:-$(nproc)}
if [ ! -d "$REF/ffmpeg/libavcodec" ]; then
  echo "build_ref: $REF not present; keeping prebuilt oracle/_ref as is" >&2
  exit 0
fi
if [ -x "$OUT/recode_ref" ] && [ -x "$OUT/recode_ref_dbg" ] && [ -x "$OUT/recode_ref_h8" ] \
   && [ -x "$OUT/recode_ref_trace" ] && [ "${FORCE:-0}" != "1" ]; then
  echo "build_ref: oracle/_ref already built" >&2
  exit 0
fi
mkdir -p "$B/ff"
cd "$B/ff"
if [ ! -f libavcodec/libavcodec.a ]; then
  "$REF/ffmpeg/configure" --disable-yasm --disable-inline-asm --disable-programs --disable-doc \
    --disable-everything --enable-decoder=h264 --enable-parser=h264 --enable-demuxer=h264 \
    --enable-demuxer=mov --enable-demuxer=matroska --enable-protocol=file \
    --disable-avdevice --disable-avfilter --disable-swscale --disable-swresample \
    --disable-postproc --disable-network --disable-zlib --disable-bzlib --disable-lzma \
    --disable-iconv --disable-xlib --disable-sdl > configure.log 2>&1
  sed -i 's/#define HAVE_ARC4RANDOM 1/#define HAVE_ARC4RANDOM 0/' config.h
  make -j"$JOBS" > make.log 2>&1
fi
cd "$B"
sed -e 's/std::function<void(T\*&)>/std::function<void(T*)>/g' \
    -e 's/(T\*& to_delete)/(T* to_delete)/g' "$REF/recode.cpp" > recode_patched.cpp
# H8: decompressor learns is_8x8 at begin_coding_type like the compressor does before its NZ bits
awk '{print} /bool begin_coding_type\(CodingType ct, int zz_index, int param0, int param1\) \{/ && !done {getline; print; print "    if (ct == PIP_SIGNIFICANCE_MAP) { BlockMeta &m8 = frames[cur_frame].meta_at(mb_coord.mb_x, mb_coord.mb_y); m8.is_8x8 = m8.is_8x8 || sub_mb_size > 32; }"; done=1}' \
    recode_patched.cpp > recode_h8.cpp
# trace variant: log every model-coded bin
python3 - "$B" <<'PY'
import re, sys
b = sys.argv[1]
s = open(b + "/recode_patched.cpp").read()
hdr = '''
#include <cstdio>
#include <cstdlib>
static FILE* avr_trace_file() { static FILE* f = getenv("AVR_TRACE") ? fopen(getenv("AVR_TRACE"), "w") : nullptr; return f; }
'''
s = s.replace('#include "framebuffer.h"', '#include "framebuffer.h"\n' + hdr, 1)
# in update_state_for_model_key: print key + symbol
s = s.replace('    auto* e = &estimators[key];\n    if (symbol) {',
  '    auto* e = &estimators[key];\n    if (FILE* tf = avr_trace_file()) { fprintf(tf, "%d %d %d %d %d %d %d\\n", (int)coding_type, (int)((const uint8_t*)std::get<0>(key) - (const uint8_t*)this), std::get<1>(key), std::get<2>(key), symbol, e->pos, e->neg); }\n    if (symbol) {', 1)
open(b + "/recode_trace.cpp", "w").write(s)
PY
INC="-I$HERE/ref_shim -I$REF -I$REF/ffmpeg -I$B/ff"
LIBS="$B/ff/libavformat/libavformat.a $B/ff/libavcodec/libavcodec.a $B/ff/libavutil/libavutil.a -lm -lpthread"
CXX="g++ -std=c++14 -w -fpermissive -include stdexcept -include cstring"
$CXX -O2 -DNDEBUG $INC recode_patched.cpp $LIBS -o "$OUT/recode_ref" &
$CXX -O2          $INC recode_patched.cpp $LIBS -o "$OUT/recode_ref_dbg" &
$CXX -O2 -DNDEBUG $INC recode_h8.cpp      $LIBS -o "$OUT/recode_ref_h8" &
$CXX -O2          $INC recode_trace.cpp   $LIBS -o "$OUT/recode_ref_trace" &
wait
# the reference's only first-party unit test (test/arithmetic_code.cpp), for the KAT in tests/
$CXX -O2 $INC "$REF/test/arithmetic_code.cpp" -o "$OUT/test_arithmetic_code" || true
ls -la "$OUT"
