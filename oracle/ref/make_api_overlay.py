"""TEST INFRASTRUCTURE.  The attach mechanism of SURVEY.md §8(b): a scripted two-hunk overlay of the reference's
src/core/api.cpp, written to the git-ignored build directory (oracle/_ref/objfull/api_overlay.cpp — the reference tree is
never edited and no reference source is committed), that adds the two name branches a maintainer would add by hand
(INTEGRATION.md):

    MakeAccelerator  (api.cpp:790-1015):   else if (name == "b200") accel = CreateB200Accelerator(std::move(prims), paramSet);
    MakeIntegrator   (api.cpp:1895-1927):  else if (IntegratorName == "b200wavefront")
                                               integrator = CreateB200WavefrontIntegrator(IntegratorParams, sampler, camera);

usage: make_api_overlay.py <reference api.cpp> <output>"""
import sys

src, out = sys.argv[1], sys.argv[2]
s = open(src).read()

inc = '#include "media/homogeneous.h"\n'
assert s.count(inc) == 1
s = s.replace(inc, inc + '#include "b200accel.h"\n#include "b200integrator.h"\n')

hunk1 = '    else\n        Warning("Accelerator \\"%s\\" unknown.", name.c_str());\n'
assert s.count(hunk1) == 1, "MakeAccelerator's unknown-name branch not found"
s = s.replace(hunk1, '    else if (name == "b200")\n        accel = CreateB200Accelerator(std::move(prims), paramSet);\n' + hunk1)

hunk2 = '    } else {\n        Error("Integrator \\"%s\\" unknown.", IntegratorName.c_str());\n'
assert s.count(hunk2) == 1, "MakeIntegrator's unknown-name branch not found"
s = s.replace(hunk2, '    } else if (IntegratorName == "b200wavefront") {\n'
                     '        integrator = CreateB200WavefrontIntegrator(IntegratorParams, sampler, camera);\n' + hunk2)
open(out, "w").write(s)
