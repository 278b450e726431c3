// stand-in: Sophus::SE3f / Sim3f live in prelude.h (force-included); the reference header is not in the image
#pragma once
