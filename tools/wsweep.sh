for w in 32 16 8 4 2; do echo "wpw $w"; MYO_WPW=$w tools/sweep.sh 8192; done
echo "auto"; tools/sweep.sh 8192 32768 65536
