"""B200-native ORB front-end: drop-in for ORB-SLAM3's ORBextractor / ORBmatcher hot path."""
