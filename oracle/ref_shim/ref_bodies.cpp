// ref_bodies.cpp — the translation unit that holds the reference's method bodies, included VERBATIM from the line ranges that
// oracle/ref_build.py copies out of /root/reference into oracle/_ref/gen/ (git-ignored build output; never committed).
// TEST INFRASTRUCTURE ONLY.  Nothing in this file is written by hand except the include list.
#include "ref_shells.hpp"

using namespace std::chrono;  // MultiCameraFrame.cpp:10

#include "gen/tools_build_Rt.inc"                 // common_utils/src/tools.cpp:136-148
#include "gen/mcf_extractFeatures.inc"            // MultiCameraFrame.cpp:203-262   extractFeaturesParallel + extractFeatureSingle
#include "gen/mcf_UndistortKeyPoints.inc"         // MultiCameraFrame.cpp:300-347
#include "gen/mcf_triangulateIntraMatches.inc"    // MultiCameraFrame.cpp:411-441   N-view DLT
#include "gen/mcf_triangulate.inc"                // MultiCameraFrame.cpp:480-507   2-view f32 DLT
#include "gen/mcf_computeRepresentativeDesc.inc"  // MultiCameraFrame.cpp:530-567
#include "gen/mcf_BruteForceMatch.inc"            // MultiCameraFrame.cpp:1024-1086
#include "gen/mcf_computeIntraMatches.inc"        // MultiCameraFrame.cpp:1100-1288 computeIntraMatches(matches, old)
#include "gen/frontend_obtainLfFeatures.inc"      // FrontEnd.cpp:213-593
