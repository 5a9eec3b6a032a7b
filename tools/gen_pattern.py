"""Regenerates include/mcs_orb_pattern.inc (a DATA table) from the reference's bit_pattern_31_
(MCSlam/src/ORBextractor.cpp:148-406). Only runnable where /root/reference is mounted."""
import re, sys
src = open('/root/reference/MCSlam/src/ORBextractor.cpp').read()
m = re.search(r'bit_pattern_31_\[256\*4\]\s*=\s*\{(.*?)\};', src, re.S)
nums = [int(x) for x in re.findall(r'-?\d+', re.sub(r'/\*.*?\*/', '', m.group(1), flags=re.S))]
assert len(nums) == 1024
print(nums)
