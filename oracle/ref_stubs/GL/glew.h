// Test-infrastructure stub (oracle/_ref build only): shadows <GL/glew.h> so that the reference's
// simulator/core/System.cpp compiles headless. There is no graphics stack on this box; nothing here is called.
#pragma once
