// Test-infrastructure stub: see ImfStub.h
#pragma once
#include "ImfStub.h"
