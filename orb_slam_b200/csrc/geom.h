// Per-level geometry of one extractor configuration (host + device view).
// All sizes derive from the frame size and the five ORBextractor.* settings
// exactly as the reference derives them (src/ORBextractor.cc:466-540 ctor,
// :1347-1352 level sizes, :932-949 cell grid, :650-653 initial quadtree nodes).
#ifndef ORB_B200_GEOM_H
#define ORB_B200_GEOM_H

#include <stddef.h>
#include <stdint.h>

#define ORB_MAX_LEVELS 16
#define ORB_EDGE 19          // EDGE_THRESHOLD, src/ORBextractor.cc:74
#define ORB_BORDER_MIN 16    // minBorderX/Y = EDGE_THRESHOLD-3, :932
#define ORB_XPAD 32          // bytes before interior column 0 in every pyramid row (19 border + 13 alignment)
#define ORB_HALF_PATCH 15    // HALF_PATCH_SIZE, :73

struct LevelGeom {
    int w, h;                // level size (un-bordered), :1348
    int pitch;               // row pitch (bytes) shared by the pyramid, blur and score planes of this level
    int prow0;               // first row of the interior inside the bordered pyramid plane (= ORB_EDGE)
    unsigned long long pyrOff;    // byte offset of the bordered plane inside one frame's pyramid block
    unsigned long long planeOff;  // byte offset of the un-bordered plane inside one frame's blur / score block
    // resize tables (device arrays indexed from these offsets); level 0 has none
    int xtabOff, ytabOff;
    // FAST cell grid, :941-949
    int maxBorderX, maxBorderY;   // w-16, h-16
    int nCols, nRows, wCell, hCell;
    int cellOff;             // index of this level's first cell in the per-frame cell arrays
    // candidates
    int candCap;             // worst-case number of NMS survivors
    int slotCap;             // per-cell survivor list capacity (worst case of the largest cell)
    unsigned long long slotOff;   // element offset of the level's first cell slot list inside one frame's slot array
    unsigned long long candOff;   // element offset inside one frame's candidate array
    // quadtree, :644-907
    int N;                   // mnFeaturesPerLevel[level]
    int nIni;                // round(W/H)
    float hX;                // W / nIni
    int qW, qH;              // maxX-minX, maxY-minY
    int depth;               // number of 2-bit digits in a path code
    int iniBits;             // bits for the initial node index
    int codeXOff, codeYOff;  // this level's column / row parts of the path codes inside the code table (quadtree.cuh)
    int maxSel;              // upper bound on retained keypoints: max(N+3, 4*nIni)
    int selOff;              // slot offset inside one frame's selection array
    // output
    float scale;             // mvScaleFactor[level]
    float patchSize;         // (int)(31*scale) as float, :1013,:1024
};

struct FrameGeom {
    int nlevels;
    int width, height;
    int iniTh, minTh;
    int totalCells;          // cells per frame over all levels
    int selTotal;            // selection slots per frame (= keypoint capacity per frame)
    unsigned long long pyrBytes;     // per frame
    unsigned long long planeBytes;   // per frame (blur block and score block each)
    unsigned long long candTotal;    // candidate elements per frame
    unsigned long long slotTotal;    // cell slot elements per frame
    int nodeCap;             // quadtree node capacity (power of two) over all levels
    int presortLevels;       // levels 0 .. presortLevels-1 are large: gathered / coded / sorted by a CTA cluster each
    int presortMaxCells;     // cells of the largest of them
    LevelGeom lv[ORB_MAX_LEVELS];
};

// one entry per destination column / row of a resized level
struct ResizeTap {
    int s0, s1;              // source index of the two taps (s1 clipped; equals s0 when the weight of tap 1 is 0 at the right edge)
    short a0, a1;            // 11-bit fixed point weights
};

#endif
