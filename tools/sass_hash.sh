#!/bin/bash
# hash of the SASS of one kernel of an object file (addresses kept, so any code change shows): sass_hash.sh obj mangled-substring
cuobjdump -sass "$1" | awk -v k="$2" '/Function : /{f = index($0, k) > 0} f' | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed 's#/\* 0x[0-9a-f]* \*/##' | md5sum
