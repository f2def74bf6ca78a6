/* Empty stand-in for <emscripten/atomic.h> (reference main.cpp:5 includes it but
 * uses no symbol from it). Test infrastructure only: lets oracle/build_ref.py
 * compile the reference's x64 path with g++ on Linux. */
#pragma once
