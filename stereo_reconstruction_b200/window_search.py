"""Host-side mirror of the reference's WindowSearch interface (same names, argument meaning and
error behaviour), forwarding to the C-ABI.  Used by the parity tests so they read like calls
into the reference:

    BlockSearch(left, right, 17, 0, 200).computeDisparityMapRight(0.9)      # BlockSearch.h:9-46
    LinearSearch(left, right).computeDisparityMap(1.0)                      # LinearSearch.h:9-20

Images are HxWx3 uint8 BGR numpy arrays (cv::Mat CV_8UC3); results are float64 maps (CV_64F).
The C++ equivalents for a build with OpenCV are in shim/.
"""
from . import capi


class BlockSearch:
    """BlockSearch(leftImage, rightImage, blockSize | blockSizes, minDisparity, maxDisparity) -- BlockSearch.cpp:6-22."""

    def __init__(self, leftImage, rightImage, blockSize, minDisparity, maxDisparity, *, cost=capi.COST_SSD_L2,
                 ctx=None):
        self.leftImage_ = leftImage      # shallow, never written (BlockSearch.h:40-41)
        self.rightImage_ = rightImage
        if isinstance(blockSize, (list, tuple)):
            # the vector overload leaves blockSize_ uninitialised in the reference (:15-22);
            # supported as blockSizes[0]
            self.blockSizes_ = list(blockSize)
            self.blockSize_ = int(blockSize[0])
        else:
            self.blockSize_ = int(blockSize)
            self.blockSizes_ = [int(blockSize)]
        self.minDisparity_ = int(minDisparity)
        self.maxDisparity_ = int(maxDisparity)
        self.cost_ = cost
        self.ctx_ = ctx or capi.default_context()
        self.maxBlockSize = self.blockSize_

    def computeDisparityMapRight(self, smoothFactor, varBlock=False, thres=19.0, *, subpixel=False, lr_check=False):
        p = capi.search_params(capi.VARIANT_RIGHT, self.blockSize_, self.minDisparity_, self.maxDisparity_,
                               smoothFactor, varBlock, thres, self.cost_, subpixel, lr_check)
        out, mb = capi.block_search(self.ctx_, self.leftImage_, self.rightImage_, p, return_max_block=True)
        self.maxBlockSize = mb
        print("max block size: ", mb)     # BlockSearch.cpp:177 (kept in the shim, not the library)
        return out

    def computeDisparityMapLeft(self, smoothFactor):
        p = capi.search_params(capi.VARIANT_LEFT, self.blockSize_, self.minDisparity_, self.maxDisparity_,
                               smoothFactor, False, 19.0, capi.COST_SSD_L2)
        return capi.block_search(self.ctx_, self.leftImage_, self.rightImage_, p)


class LinearSearch:
    """LinearSearch(leftImage, rightImage) -- LinearSearch.h:9-20."""

    def __init__(self, leftImage, rightImage, *, ctx=None):
        self.leftImage_ = leftImage
        self.rightImage_ = rightImage
        self.ctx_ = ctx or capi.default_context()

    def computeDisparityMap(self, smoothFactor):
        print(self.leftImage_.shape[0])   # LinearSearch.cpp:16-17
        print(self.rightImage_.shape[0])
        p = capi.search_params(capi.VARIANT_LINEAR, 1, 0, 200, smoothFactor)
        return capi.block_search(self.ctx_, self.leftImage_, self.rightImage_, p)
