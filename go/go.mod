module github.com/LynnColeArt/guda-b200/go

go 1.21
