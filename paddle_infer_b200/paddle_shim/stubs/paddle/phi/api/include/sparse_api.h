// Stand-in for the build-generated sparse_api.h (nothing from it is used here).
#pragma once
