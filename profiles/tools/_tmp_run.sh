#!/bin/bash
cd /root/repo
timeout 200 python profiles/tools/tile_prof.py
