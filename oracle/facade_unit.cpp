// TEST INFRASTRUCTURE. The reference's OWN unit tests of the scene path, compiled UNCHANGED against the product's hg::Scene facade:
// `make -C oracle facade_unit` cuts the test functions out of /root/reference/harfang/tests/engine/scene.cpp (by line range, into
// oracle/_ref/facade_unit_extract.inc — generated, never committed) and builds them with the reference's test framework (acutest) against
// harfang3d_b200/csrc/facade/include + libhg_facade.so:
//   test_ComponentGarbageCollection (tests/engine/scene.cpp:27-33), test_WalkHierarchy (:55-81), test_DisableObjectNodes (:115-154)
// These three run on the host alone (no device call is reached). Run by tests/test_facade.py::test_reference_unit_tests_against_facade.
#include "acutest.h"

#undef near
#undef far

#include "engine/scene.h"

using namespace hg;

#include "facade_unit_extract.inc"

TEST_LIST = {
	{"engine.scene.ComponentGarbageCollection", test_ComponentGarbageCollection},
	{"engine.scene.WalkHierarchy", test_WalkHierarchy},
	{"engine.scene.DisableObjectNodes", test_DisableObjectNodes},
	{NULL, NULL},
};
