// oracle/ref_unicode_probe.cpp -- TEST INFRASTRUCTURE.  Exposes the reference's own codepoint classifier (unicode.h:360-403,
// the tables the BPE pre-tokenizer consults) through a C symbol, so tests/test_tokenizer_cpu.py can compare the generated
// table in pipeinfer_b200/csrc/unicode_classes.inc with it over all of Unicode.  Compiled against the header where it lies
// under /root/reference (oracle/Makefile -> oracle/_ref/libunicode_probe.so); nothing of it is copied.
#include <cstdint>
#include <string>

#include "unicode.h"

extern "C" int ref_codepoint_type(uint32_t cp) { return codepoint_type(cp); }
extern "C" int ref_byte_symbol(int byte, char * out4) {
    const std::string s = bytes_to_unicode_bpe((uint8_t) byte);
    for (size_t i = 0; i < s.size(); i++) out4[i] = s[i];
    return (int) s.size();
}
